import sys
sys.path.insert(0,'opencl-raytracer_b200')
from pyrt import lib, scenes
objs, lights = scenes.synthetic_scene(100_000, 4096)
for th in (0, 0, 1):
    info, nodes, leaf = lib.host_build_bvh(objs, lights, threads=th)
    print("threads", th, "build_ms", round(info.build_ms, 1), file=sys.stderr)
