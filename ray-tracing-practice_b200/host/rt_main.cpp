// rt_main.cpp -- drop-in for the reference executable (src/main.cu:572-606): `rt_main [--gpu|--default] < config`.
// `--cpu` is refused: this build has no CPU render path (the reference's --cpu lives on as oracle/_ref).
// Extensions (all optional, after the mode): --png | --ppm choose the saver the reference ships but never wires up;
// --spheres N [--seed S] appends the seeded N-sphere field of BASELINE.json's configs 2-4; --bvh-gpu builds the BVH on
// the device instead of on the host; --gpus N renders the frames of the camera path on N devices
// (--split tiles: all N devices render every frame together, tiles gathered over NCCL); --parity renders with the reference's
// own random stream and arithmetic (frames byte-identical to the reference program's --cpu run); --no-pipeline writes each
// file before the next frame starts, like the reference.
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>

#include "scene_host.h"

int main(int argc, char** argv) {
    using namespace rtb;
    int spheres = 0;
    DriverOpts d;
    unsigned seed = 1234;
    bool want_default = false;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        if (a == "--default") want_default = true;
        else if (a == "--gpu") {}
        else if (a == "--cpu") { std::cerr << "rt_main: no CPU render path in this build (use the reference's --cpu)\n"; return 2; }
        else if (a == "--png") d.format = 1;
        else if (a == "--ppm") d.format = 2;
        else if (a == "--parity") d.parity = true;
        else if (a == "--no-pipeline") d.pipeline = false;
        else if (a == "--split" && i + 1 < argc) d.split_tiles = std::string(argv[++i]) == "tiles";
        else if (a == "--spheres" && i + 1 < argc) spheres = std::atoi(argv[++i]);
        else if (a == "--bvh-gpu") setenv("RTB_BVH_BUILDER", "gpu", 1);  // device BVH build (rt_scene_create_ex, RT_BVH_GPU_LBVH)
        else if (a == "--gpus" && i + 1 < argc) d.num_gpus = std::atoi(argv[++i]);  // frames n % gpus == g render on device g
        else if (a == "--seed" && i + 1 < argc) seed = (unsigned)std::strtoul(argv[++i], nullptr, 10);
    }
    if (want_default) { std::cout << default_config_text(); return 0; }
    SceneParams params = read_scene_params(std::cin);
    HostScene scene;
    std::string err;
    if (!create_scene(params, scene, true, &err)) { std::cerr << err << std::endl; return 1; }
    if (!err.empty()) std::cerr << err << std::endl;
    if (spheres > 0) make_synthetic_spheres(scene, spheres, seed);
    return gpu_render(params, scene, d, std::cout);
}
