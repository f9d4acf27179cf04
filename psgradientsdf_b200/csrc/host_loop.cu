// Host-side outer loop: the reference's alternatingOptimize control flow, log text and artefact files, driving the
// CUDA blocks through the same C entry points an external host would use.
//
// Mirrors PsOptimizer::alternatingOptimize (cpp/include/ps_optimizer/PsOptimizer.cpp:239-428) and
// LedOptimizer::alternatingOptimize (LedOptimizer.cpp:279-478): block order (SH: albedo, light, distance, pose;
// LED: light, albedo, distance, pose), one-off rescaling of the regulariser weights by E / E_reg, convergence and
// divergence tests on the total energy after the last enabled block, upsampling at iteration 5, optimizer_doc.txt,
// and the every-third-iteration dumps (Optimizer::savePoses / save_pointcloud / extract_mesh, OptimizerAux.cpp:278-362, 456-511,
// 580-599; the meshes come from mesh.cu).
#include <chrono>
#include <cmath>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>

#include "psg_host.h"

namespace {

struct Energies { float E = 0.f, En = 0.f, El = 0.f, Er = 0.f; };

int ps_energy(psg_optimizer* o, float* E) {
    double e = 0.0;
    const int st = psg_opt_energy(o, &e, nullptr);
    *E = (float)e;
    return st;
}

// Optimizer::getTotalEnergy (OptimizerAux.cpp:259-269)
float total_energy(psg_optimizer* o, const Energies& e, std::ofstream& file, bool echo) {
    float reg[3];
    psg_opt_get_reg(o, reg);
    const float E_total = e.E + reg[1] * e.En + reg[2] * e.El + reg[0] * e.Er;
    std::ostringstream line;
    line << "PS energy: " << e.E << "\t normal reg energy: " << reg[1] * e.En << "\t laplacian reg energy: " << reg[2] * e.El
         << "\t rho reg energy: " << reg[0] * e.Er << "\t total energy: " << E_total;
    if (echo) std::cout << line.str() << std::endl;
    if (file.is_open()) file << line.str() << std::endl;
    return E_total;
}

// Eigen::Quaternion<float>(Matrix3f) (third/eigen/Eigen/src/Geometry/Quaternion.h, quaternionbase_assign_impl<3,3>)
void quat_from_matrix(const float* P /* 4x4 row-major */, float q[4] /* x y z w */) {
    auto m = [&](int r, int c) { return P[4 * r + c]; };
    float t = m(0, 0) + m(1, 1) + m(2, 2);
    if (t > 0.f) {
        t = std::sqrt(t + 1.f);
        q[3] = 0.5f * t;
        t = 0.5f / t;
        q[0] = (m(2, 1) - m(1, 2)) * t;
        q[1] = (m(0, 2) - m(2, 0)) * t;
        q[2] = (m(1, 0) - m(0, 1)) * t;
    } else {
        int i = 0;
        if (m(1, 1) > m(0, 0)) i = 1;
        if (m(2, 2) > m(i, i)) i = 2;
        const int j = (i + 1) % 3, k = (j + 1) % 3;
        t = std::sqrt(m(i, i) - m(j, j) - m(k, k) + 1.f);
        q[i] = 0.5f * t;
        t = 0.5f / t;
        q[3] = (m(k, j) - m(j, k)) * t;
        q[j] = (m(j, i) + m(i, j)) * t;
        q[k] = (m(k, i) + m(i, k)) * t;
    }
}

// Optimizer::savePoses (OptimizerAux.cpp:580-599)
bool save_poses(psg_optimizer* o, const std::string& path) {
    const std::vector<std::string>& stamps = o->key_stamps;
    std::ofstream f(path.c_str());
    if (!f.is_open()) return false;
    for (int i = 0; i < o->F; ++i) {
        const float* P = o->poses.data() + 16 * i;
        float q[4];
        quat_from_matrix(P, q);
        const std::string stamp = i < (int)stamps.size() ? stamps[i] : std::to_string(o->frame_idx[i]);
        f << stamp << " " << P[3] << " " << P[7] << " " << P[11] << " " << q[0] << " " << q[1] << " " << q[2] << " " << q[3] << "\n";
    }
    return true;
}

// Optimizer::save_pointcloud (OptimizerAux.cpp:456-511): grid-relative position vs * idx - d * g_hat, normal, colour
bool save_pointcloud(psg_optimizer* o, const std::string& path) {
    psg_volume* v = o->vol;
    // only the surface voxels leave the device (S x 32 B), not the volume (N x 32 B: 4 GiB at 512^3)
    std::vector<float> rec((size_t)(o->S > 0 ? o->S : 1) * 8);
    if (psg_opt_get_surface_voxels(o, rec.data()) != PSG_OK) return false;
    std::vector<int> lin((size_t)(o->S > 0 ? o->S : 1));
    if (psg_opt_get_surface(o, lin.data(), nullptr) != PSG_OK) return false;
    const double thr = std::sqrt(3.0) * (double)o->vs;
    std::ostringstream body;
    size_t count = 0;
    const int W0 = v->dims[0], WH = v->dims[0] * v->dims[1];
    for (int s = 0; s < o->S; ++s) {
        const float* r = rec.data() + (size_t)s * 8;
        if (!(std::fabs((double)r[0]) < thr)) continue;
        int rest = lin[s];
        const int k = rest / WH; rest -= k * WH;
        const int j = rest / W0; rest -= j * W0;
        const int i = rest;
        float g[3] = { r[1], r[2], r[3] };
        const float z = g[0] * g[0] + (g[1] * g[1] + g[2] * g[2]);
        if (z > 0.f) { const float sq = std::sqrt(z); g[0] /= sq; g[1] /= sq; g[2] /= sq; }
        const float p[3] = { o->vs * (float)i - r[0] * g[0], o->vs * (float)j - r[0] * g[1], o->vs * (float)k - r[0] * g[2] };
        body << p[0] << " " << p[1] << " " << p[2] << " " << g[0] << " " << g[1] << " " << g[2] << " " << int(255 * r[5]) << " "
             << int(255 * r[6]) << " " << int(255 * r[7]) << std::endl;
        ++count;
    }
    std::ofstream ply(path.c_str());
    if (!ply.is_open()) return false;
    ply << "ply" << std::endl << "format ascii 1.0" << std::endl << "element vertex " << count << std::endl;
    for (const char* pn : {"x", "y", "z", "nx", "ny", "nz"}) ply << "property float " << pn << std::endl;
    for (const char* pn : {"red", "green", "blue"}) ply << "property uchar " << pn << std::endl;
    ply << "end_header" << std::endl << body.str();
    return true;
}

// Optimizer::saveSDF (OptimizerAux.cpp:513-577): cropped box of -dist values; only the box leaves the device (mesh.cu)
bool save_sdf(psg_optimizer* o, const std::string& path) {
    try { bool any = false; return psg::write_sdf_file(o->vol, o->vs, path.c_str(), &any) && any; }
    catch (const psg::CudaError&) { return false; }
}

// Optimizer::extract_mesh (OptimizerAux.cpp:278-362): <name>_mesh.ply, marching cubes on the device
bool save_mesh(psg_optimizer* o, const std::string& path) {
    try { long long nt = 0; return psg::write_mesh_file(o->vol, o->vs, path.c_str(), 0, &nt) && nt > 0; }
    catch (const psg::CudaError&) { return false; }
}

}  // namespace

extern "C" int psg_opt_save_pointcloud(psg_optimizer* o, const char* filename) {
    if (!o || !filename) { psg::set_error("psg_opt_save_pointcloud: null argument"); return PSG_ERR_INVALID; }
    if (!o->inited) { psg::set_error("psg_opt_save_pointcloud: psg_opt_init has not run"); return PSG_ERR_STATE; }
    if (!save_pointcloud(o, filename)) { psg::set_error("psg_opt_save_pointcloud: cannot write %s", filename); return PSG_ERR_INVALID; }
    return PSG_OK;
}

extern "C" int psg_opt_alternating_optimize(psg_optimizer* o, const psg_loop_config* cfg, const char* save_path,
                                            psg_loop_result* res) {
    if (!o || !cfg) { psg::set_error("psg_opt_alternating_optimize: null argument"); return PSG_ERR_INVALID; }
    if (!o->inited) { psg::set_error("psg_opt_init must be called first"); return PSG_ERR_STATE; }
    const bool led = o->set.model == PSG_MODEL_LED;
    const std::string out = save_path ? save_path : "";
    const bool files = cfg->write_artifacts != 0 && save_path != nullptr && o->rank == 0;   // several GPUs: rank 0 writes
    const bool echo = cfg->verbose != 0;
    psg_loop_result R;
    memset(&R, 0, sizeof(R));
    std::ofstream file;
    if (files) file.open((out + "optimizer_doc.txt").c_str());

    float reg[3];
    psg_opt_get_reg(o, reg);
    bool normal_reg = reg[1] != 0.f, albedo_reg = reg[0] != 0.f, laplacian_reg = reg[2] != 0.f;
    const float damping = cfg->damping;
    if (file.is_open())
        file << "albation study settings: \t" << "light: " << (cfg->light != 0) << "\t" << "albedo: " << (cfg->albedo != 0) << "\t"
             << "distance: " << (cfg->distance != 0) << "\t" << "pose: " << (cfg->pose != 0) << "\n"
             << "num of key frame: " << o->F << " \n total voxels: " << o->vol->N << "\n";

#define PSG_RUN(call) do { const int st__ = (call); if (st__ != PSG_OK) return st__; } while (0)
    PSG_RUN(psg_opt_init_albedo(o));
    Energies e;
    PSG_RUN(ps_energy(o, &e.E));
    double dn = 0, dl = 0, dr = 0;
    if (normal_reg) {
        PSG_RUN(psg_opt_reg_energies(o, &dn, nullptr, nullptr));
        e.En = (float)dn;
        reg[1] *= e.E / e.En;   // settings_->reg_weight_n *= E / E_n (PsOptimizer.cpp:277)
    }
    if (albedo_reg) { PSG_RUN(psg_opt_reg_energies(o, nullptr, nullptr, &dr)); e.Er = (float)dr; }
    if (laplacian_reg) {
        PSG_RUN(psg_opt_reg_energies(o, nullptr, &dl, nullptr));
        e.El = (float)dl;
        reg[2] *= e.E / e.El;
        if (cfg->upsample) laplacian_reg = false;
    }
    psg_opt_set_reg(o, reg);
    float E_total = total_energy(o, e, file, echo);
    std::vector<float> E_vec(1, E_total);

    // The reference evaluates getPSEnergy() after every block.  Here every block's own sweep already yields the energy of
    // the state it started from, and psg_opt_iteration queues the blocks with three host waits per pass; the log lines of an
    // iteration are written once it has returned, in the reference's order and with the values the reference would print.
    const char* labels[4];
    labels[PSG_BLOCK_ALBEDO] = "albedo"; labels[PSG_BLOCK_LIGHT] = "light"; labels[PSG_BLOCK_DIST] = "distance"; labels[PSG_BLOCK_POSE] = "pose";
    const int order_sh[4] = { PSG_BLOCK_ALBEDO, PSG_BLOCK_LIGHT, PSG_BLOCK_DIST, PSG_BLOCK_POSE };
    const int order_led[4] = { PSG_BLOCK_LIGHT, PSG_BLOCK_ALBEDO, PSG_BLOCK_DIST, PSG_BLOCK_POSE };
    const int* order = led ? order_led : order_sh;
    auto run_iteration = [&]() -> int {
        psg_iteration_config ic;
        memset(&ic, 0, sizeof(ic));
        ic.light = cfg->light; ic.albedo = cfg->albedo; ic.distance = cfg->distance; ic.pose = cfg->pose;
        ic.damping = damping;
        ic.reg_flags = (normal_reg ? PSG_REG_NORMAL : 0) | (laplacian_reg ? PSG_REG_LAPLACIAN : 0) | (albedo_reg ? PSG_REG_ALBEDO : 0);
        psg_iteration_info info;
        const int st = psg_opt_iteration(o, &ic, &info);
        if (st != PSG_OK) return st;
        for (int k = 0; k < 4; ++k) {
            const int block = order[k];
            if (!info.ran[block]) continue;
            e.E = (float)info.energy_after[block];
            if (block == PSG_BLOCK_DIST) {
                if (normal_reg) e.En = (float)info.reg_normal_after_dist;
                if (laplacian_reg) e.El = (float)info.reg_laplacian_after_dist;
            }
            if (block == PSG_BLOCK_ALBEDO && albedo_reg) e.Er = (float)info.reg_albedo_after_albedo;   // PsOptimizer.cpp:311
            if (echo) std::cout << "===> [" << R.iterations << "]: after " << labels[block] << " optimization: ";
            if (file.is_open()) file << "===> [" << R.iterations << "]: after " << labels[block] << " optimization: \n";
            E_total = total_energy(o, e, file, echo);
        }
        return PSG_OK;
    };

    double ms_sum = 0.0;
    int iter = 0;
    int verdict = 0;   // 1 converged, 2 diverged
    while (iter < cfg->max_it) {
        R.iterations = iter;
        const auto t0 = std::chrono::steady_clock::now();
        PSG_RUN(run_iteration());
        const auto t1 = std::chrono::steady_clock::now();
        const double ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
        ms_sum += ms;
        if (iter < 64) R.iteration_ms[iter] = (float)ms;

        E_vec.push_back(E_total);
        const float prev = E_vec[E_vec.size() - 2];
        const float rel_diff = std::fabs(prev - E_total) / prev;
        if (echo) std::cout << "===> [" << iter << "]: relative diff " << rel_diff << std::endl;
        if (file.is_open()) file << "===> [" << iter << "]: relative diff " << rel_diff << "\n";
        R.last_rel_diff = rel_diff;

        if (rel_diff < cfg->conv_threshold) {
            if (file.is_open()) file << "===> [" << iter << "]: converged! \n";
            verdict = 1;
        } else if (prev < E_total) {
            if (file.is_open() && !led) file << "===> [" << iter << "]: diverged!\n";   // the LED path logs this to stdout only
            verdict = 2;
        }
        if (verdict) {
            PSG_RUN(psg_opt_sync_volume(o));   // several GPUs, sharded: every rank ends with the whole reconstruction
            if (files) {
                save_pointcloud(o, out + "final_refined_pointcloud.ply");
                save_mesh(o, out + "final_refined_mesh.ply");
                if (led) save_sdf(o, out + "refined_sdf.sdf");
            }
            ++iter;
            break;
        }

        if (iter == 5 && cfg->upsample) {
            psg_opt_get_reg(o, reg);
            if (reg[2] == 0.f) reg[2] = 1.f;
            psg_opt_set_reg(o, reg);
            laplacian_reg = true;
            PSG_RUN(psg_opt_upsample(o));
            if (files) {
                save_pointcloud(o, out + "upsample_after_" + std::to_string(iter) + "_pointcloud.ply");
                save_mesh(o, out + "upsample_after_" + std::to_string(iter) + "_mesh.ply");
            }
            if (file.is_open()) file << "===> [" << iter << "]: after pose optimization: \n";
            PSG_RUN(psg_opt_reg_energies(o, nullptr, &dl, nullptr));
            e.El = (float)dl;
            psg_opt_get_reg(o, reg);
            reg[2] *= e.E / e.El;
            psg_opt_set_reg(o, reg);
            E_total = total_energy(o, e, file, echo);
            E_vec.push_back(E_total);
        }
        // SH: iter > 15 (PsOptimizer.cpp:411); LED: iter == 15 (LedOptimizer.cpp:461)
        if (cfg->upsample && (led ? iter == 15 : iter > 15)) {
            psg_opt_get_reg(o, reg);
            reg[2] = 0.f;
            psg_opt_set_reg(o, reg);
        }
        ++iter;
        if (iter % 3 == 0 && cfg->write_artifacts != 0 && save_path != nullptr) PSG_RUN(psg_opt_sync_volume(o));   // collective: all ranks, not only the writer
        if (iter % 3 == 0 && files) {
            save_poses(o, out + "after_poses_opt_" + std::to_string(iter) + ".txt");
            save_pointcloud(o, out + "after_iter_" + std::to_string(iter) + "_pointcloud.ply");
            save_mesh(o, out + "after_iter_" + std::to_string(iter) + "_mesh.ply");
        }
    }
    PSG_RUN(psg_opt_sync_volume(o));
#undef PSG_RUN
    R.iterations = iter;
    R.converged = verdict == 1;
    R.diverged = verdict == 2;
    R.E = e.E; R.E_normal = e.En; R.E_laplacian = e.El; R.E_albedo = e.Er; R.E_total = E_total;
    R.ms_per_iteration = iter > 0 ? ms_sum / iter : 0.0;
    if (res) *res = R;
    return PSG_OK;
}
