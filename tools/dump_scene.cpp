// Fixture generator: loads a reference *.txt scene with the reference's OWN
// SceneLoader.cpp / ObjectData.cpp (compiled where they lie, against the compat
// glm stand-in) and writes the resulting ObjectData / Light arrays in the C-ABI
// layout of include/rt_b200.h.  Run by tools/make_fixtures.sh in the build
// container; outputs are committed under tests/golden/scenes/.
//
// File layout (little endian): "RTSC" u32 version=1 u32 n_objs u32 n_lights,
// n_objs x rt_object (184 B), n_lights x rt_light (52 B).
#include <cstdio>
#include <cstring>
#include <iostream>
#include <vector>

#include "Light.hpp"
#include "ObjectData.hpp"
#include "SceneLoader.hpp"
#include "rt_b200.h"

int main(int argc, char** argv) {
    if (argc != 3) {
        std::fprintf(stderr, "usage: %s scene.txt out.scene.bin\n", argv[0]);
        return 2;
    }
    std::vector<ObjectData> objects;
    std::vector<Light> lights;
    try {
        SceneLoader loader;
        loader.Load(argv[1], objects, lights);
    } catch (const std::exception& e) {
        std::fprintf(stderr, "%s\n", e.what());
        return 1;
    }
    FILE* f = std::fopen(argv[2], "wb");
    if (!f) return 1;
    const uint32_t hdr[3] = {1u, (uint32_t)objects.size(), (uint32_t)lights.size()};
    std::fwrite("RTSC", 1, 4, f);
    std::fwrite(hdr, 4, 3, f);
    for (const ObjectData& o : objects) {
        rt_object r;
        std::memset(&r, 0, sizeof r);
        std::memcpy(r.mat.ambient, &o.mat.ambient.x, 12);
        std::memcpy(r.mat.diffuse, &o.mat.diffuse.x, 12);
        std::memcpy(r.mat.specular, &o.mat.specular.x, 12);
        r.mat.absorption = o.mat.absorption;
        r.mat.reflection = o.mat.reflection;
        r.mat.transparency = o.mat.transparency;
        r.mat.shininess = o.mat.shininess;
        std::memcpy(r.mv, glm::value_ptr(o.mv), 64);
        std::memcpy(r.mvInverse, glm::value_ptr(o.mvInverse), 64);
        r.type = (uint32_t)o.type;
        std::fwrite(&r, sizeof r, 1, f);
    }
    for (const Light& l : lights) {
        rt_light r;
        std::memcpy(r.ambient, &l.ambient.x, 12);
        std::memcpy(r.diffuse, &l.diffuse.x, 12);
        std::memcpy(r.specular, &l.specular.x, 12);
        std::memcpy(r.position, &l.lightPosition.x, 16);
        std::fwrite(&r, sizeof r, 1, f);
    }
    std::fclose(f);
    std::printf("%s: %zu objects, %zu lights -> %s\n", argv[1], objects.size(), lights.size(), argv[2]);
    return 0;
}
