// A C++ host driving the hot path through the C ABI alone (include/psg.h, libpsg_b200.so) in the order the reference's
// main_ps.cpp drives its classes:
//
//   new VolumetricGradSdf(...), set_zmin / set_zmax                      main_ps.cpp:183-186   psg_volume_create
//   new PsOptimizer / LedOptimizer(tSDF, voxel_size, K, output, ...)     main_ps.cpp:193-202   psg_opt_create
//   per frame: increase_counter(); update(color, depth, K, pose, NEst)   main_ps.cpp:207-219   psg_volume_integrate
//   extract_mesh("init_mesh.ply"), extract_pc("init_pointcloud.ply")     main_ps.cpp:283-303   psg_volume_extract_mesh / _pc
//   setImages, setKeyframes, setPoses, init()                            main_ps.cpp:323-329   psg_opt_set_keyframes, psg_opt_init
//   alternatingOptimize(light, albedo, distance, pose)                   main_ps.cpp:330       psg_opt_alternating_optimize
//
// Frames come from one binary file (tools/write_host_input.py writes it from the synthetic scene; a real host would hand over
// what its image loader and normal estimator produced):
//   int32  F, H, W, G, model, max_it          float32 voxel_size, T, zmin, zmax, conv_threshold, damping, reg_n, reg_l
//   float32 K[9], n_sq_inv[H*W]
//   per frame: R[9], t[3] (pose.rotationMatrix(), pose.translation()), pose_init[16], color[H*W*3] (BGR), depth, nx, ny, nz [H*W]
//
//   g++ -std=c++17 -O2 -Iinclude examples/voxelps_host.cpp -Lpsgradientsdf_b200 -lpsg_b200 -Wl,-rpath,'$ORIGIN' -o psgradientsdf_b200/voxelps_host
//   psgradientsdf_b200/voxelps_host input.bin out_dir/
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "psg.h"

static void check(int status, const char* what) {
    if (status != PSG_OK) {
        std::fprintf(stderr, "%s failed (%d): %s\n", what, status, psg_last_error());
        std::exit(1);   // main_ps.cpp returns 1 on its fatal errors
    }
}

template <class T>
static std::vector<T> read_n(std::FILE* f, size_t n) {
    std::vector<T> v(n);
    if (std::fread(v.data(), sizeof(T), n, f) != n) { std::fprintf(stderr, "input file too short\n"); std::exit(1); }
    return v;
}

int main(int argc, char** argv) {
    if (argc < 3) { std::fprintf(stderr, "usage: %s input.bin out_dir/\n", argv[0]); return 1; }
    std::FILE* f = std::fopen(argv[1], "rb");
    if (!f) { std::fprintf(stderr, "cannot open %s\n", argv[1]); return 1; }
    const std::string out = argv[2];
    const std::vector<int> hi = read_n<int>(f, 6);
    const std::vector<float> hf = read_n<float>(f, 8);
    const int F = hi[0], H = hi[1], W = hi[2], G = hi[3], model = hi[4], max_it = hi[5];
    const float vs = hf[0], T = hf[1], zmin = hf[2], zmax = hf[3], conv = hf[4], damping = hf[5], reg_n = hf[6], reg_l = hf[7];
    const std::vector<float> K = read_n<float>(f, 9);
    const size_t px = (size_t)H * W;
    const std::vector<float> n_sq_inv = read_n<float>(f, px);

    const int dims[3] = { G, G, G };
    const float shift[3] = { 0.f, 0.f, 0.f };
    psg_volume* vol = nullptr;
    check(psg_volume_create(&vol, dims, vs, shift, T, zmin, zmax, /*max_frames=*/64, /*device=*/0), "psg_volume_create");
    psg_settings set{};
    set.model = model; set.loss = PSG_LOSS_CAUCHY; set.lambda = 0.2f;
    set.reg_weight_rho = 0.f; set.reg_weight_n = reg_n; set.reg_weight_l = reg_l;
    psg_optimizer* opt = nullptr;
    check(psg_opt_create(&opt, vol, vs, K.data(), &set), "psg_opt_create");   // created before the first update, as main does

    std::vector<float> images((size_t)F * px * 3), poses((size_t)F * 16);
    std::vector<int> frame_idx(F);
    for (int i = 0; i < F; ++i) {
        const std::vector<float> R = read_n<float>(f, 9), t = read_n<float>(f, 3), pose_init = read_n<float>(f, 16);
        const std::vector<float> color = read_n<float>(f, px * 3), depth = read_n<float>(f, px);
        const std::vector<float> nx = read_n<float>(f, px), ny = read_n<float>(f, px), nz = read_n<float>(f, px);
        check(psg_volume_integrate(vol, color.data(), depth.data(), nx.data(), ny.data(), nz.data(), n_sq_inv.data(), H, W,
                                   K.data(), R.data(), t.data(), /*increase_counter=*/i > 0), "psg_volume_integrate");
        std::copy(color.begin(), color.end(), images.begin() + (size_t)i * px * 3);   // every frame is a keyframe here
        std::copy(pose_init.begin(), pose_init.end(), poses.begin() + (size_t)i * 16);
        frame_idx[i] = i;
    }
    std::fclose(f);

    long long n_tri = 0, n_pts = 0;
    check(psg_volume_extract_mesh(vol, (out + "init_mesh.ply").c_str(), 0, &n_tri), "psg_volume_extract_mesh");
    check(psg_volume_extract_pc(vol, (out + "init_pointcloud.ply").c_str(), &n_pts), "psg_volume_extract_pc");

    check(psg_opt_set_keyframes(opt, F, frame_idx.data(), poses.data(), images.data(), H, W), "psg_opt_set_keyframes");
    check(psg_opt_init(opt), "psg_opt_init");
    int S = 0;
    check(psg_opt_num_surface(opt, &S), "psg_opt_num_surface");

    psg_loop_config cfg{};
    cfg.light = cfg.albedo = cfg.distance = cfg.pose = 1;
    cfg.max_it = max_it; cfg.conv_threshold = conv; cfg.damping = damping; cfg.upsample = 0;
    cfg.write_artifacts = 1; cfg.verbose = 0;
    psg_loop_result res{};
    check(psg_opt_alternating_optimize(opt, &cfg, out.c_str(), &res), "psg_opt_alternating_optimize");
    std::printf("{\"surface_voxels\": %d, \"init_mesh_triangles\": %lld, \"init_points\": %lld, \"iterations\": %d, \"converged\": %d, "
                "\"diverged\": %d, \"E_total\": %.9g, \"kernel_launches\": %lld}\n",
                S, n_tri, n_pts, res.iterations, res.converged, res.diverged, res.E_total, psg_kernel_launches());
    check(psg_opt_destroy(opt), "psg_opt_destroy");
    check(psg_volume_destroy(vol), "psg_volume_destroy");
    return 0;
}
