// trajopt_node.cpp — the C ABI used from C++, no Python in the process: the tick of the reference's C++ driver
// (TrajoptNode::stateCallback -> TrajoptSolver::shiftTrajectory + runTrajoptIteration,
// indy7-sdk/src/trajopt/src/trajopt_node.cu:105-139, include/trajopt_solver.cuh:97-175) written against
// include/indy7_b200.h.  The class keeps the reference's method names and its bookkeeping (last_update_time_,
// traj_offset_, the goal window copied BEFORE the offset moves, x_N = x_{N-1} / u_{N-1} = 0 behind a shift, x_0 = the
// measured state); the solve is i7_solve instead of sqpSolvePcg.  Host vectors only: the library stages them itself.
//
//   g++ -O2 -std=c++17 -Iinclude examples/trajopt_node.cpp -Lindy7_mpc_b200 -lindy7_b200 -Wl,-rpath,$PWD/indy7_mpc_b200
//   ./a.out model.bin goal_traj.bin N ticks        (both files: raw doubles; written by tests/test_gpu_cpp_host.py)
#include "indy7_b200.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>
#include <vector>

template <typename T> static std::vector<T> read_raw(const char* path) {
    FILE* f = std::fopen(path, "rb");
    if (!f) throw std::runtime_error(std::string("cannot open ") + path);
    std::fseek(f, 0, SEEK_END);
    const long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    std::vector<T> v(n / sizeof(T));
    if (std::fread(v.data(), sizeof(T), v.size(), f) != v.size()) { std::fclose(f); throw std::runtime_error("short read"); }
    std::fclose(f);
    return v;
}

class TrajoptSolver {      // public surface of TrajoptSolver<T, 12, 6, KnotPoints, ...> as trajopt_node.cu uses it
public:
    TrajoptSolver(const i7_model& model, const std::vector<double>& goal_ee_traj, int knot_points, double timestep)
        : N_(knot_points), dt_(timestep), goal_(goal_ee_traj), goal_len_((unsigned)(goal_ee_traj.size() / 6)) {
        i7_config cfg;
        i7_default_config(&cfg);
        cfg.N = N_; cfg.batch = 1; cfg.dt = dt_; cfg.model = model;
        if (i7_create(&cfg, &h_) != I7_OK) throw std::runtime_error(std::string("i7_create: ") + i7_last_error(nullptr));
        xu_.assign(18 * N_ - 6, 0.0);
        ee_.assign(6 * N_, 0.0);
        stats_.assign(I7_STATS_LEN, 0.0);
        copyWindow();
    }
    ~TrajoptSolver() { i7_destroy(h_); }

    void shiftTrajectory(const std::vector<double>& current_state, double current_time) {      // trajopt_solver.cuh:97-175
        if (last_update_time_ == 0) { last_update_time_ = current_time; return; }
        const double since = current_time - last_update_time_;
        int num_shifts = (int)(since / dt_);
        const double since_shift = since - num_shifts * dt_;
        num_shifts = std::min(num_shifts, (int)(goal_len_ - traj_offset_));
        if (num_shifts > 0) {
            copyWindow();                                         // at the offset BEFORE it moves, like the reference
            const int steps = std::min(num_shifts, N_ - 1);
            traj_offset_ += (unsigned)num_shifts;
            std::vector<double> shifted(xu_.size());
            check(i7_shift(h_, 1, steps, 1, xu_.data(), nullptr, nullptr, shifted.data(), nullptr, nullptr));
            xu_.swap(shifted);
            last_update_time_ = current_time - since_shift;
        }
        std::copy(current_state.begin(), current_state.begin() + 12, xu_.begin());      // x_0 = the measured state
    }

    std::string runTrajoptIteration() {
        std::vector<double> out(xu_.size());
        check(i7_solve(h_, 1, xu_.data(), xu_.data(), ee_.data(), 6, out.data(), stats_.data(), nullptr));
        xu_.swap(out);
        return "SQP iterations: " + std::to_string((int)stats_[0]) + ", Exit status: " + (stats_[3] == 0 ? "Success" : "Failure");
    }

    const std::vector<double>& getOptimizedTrajectory() const { return xu_; }
    const std::vector<double>& stats() const { return stats_; }
    bool isTrajectoryComplete() const { return traj_offset_ >= goal_len_; }
    unsigned getTrajectoryOffset() const { return traj_offset_; }
    int numKnotPoints() const { return N_; }

private:
    void check(i7_status rc) { if (rc != I7_OK) throw std::runtime_error(i7_last_error(h_)); }
    void copyWindow() {      // the next N goal points, the final position repeated past the end (trajopt_solver.cuh:120-148)
        check(i7_goal_window(h_, 1, goal_.data(), (int)goal_len_, 6, nullptr, (int)traj_offset_, 0, 6, ee_.data(), nullptr));
    }
    int N_;
    double dt_;
    std::vector<double> goal_, xu_, ee_, stats_;
    unsigned goal_len_, traj_offset_ = 0;
    double last_update_time_ = 0;
    i7_handle* h_ = nullptr;
};

int main(int argc, char** argv) {
    if (argc < 5) { std::fprintf(stderr, "usage: %s model.bin goal_traj.bin N ticks\n", argv[0]); return 2; }
    try {
        const std::vector<double> mraw = read_raw<double>(argv[1]);
        if (mraw.size() * sizeof(double) != sizeof(i7_model)) throw std::runtime_error("model file does not hold an i7_model");
        i7_model model;
        std::copy(mraw.begin(), mraw.end(), reinterpret_cast<double*>(&model));
        const std::vector<double> goal = read_raw<double>(argv[2]);
        const int N = std::atoi(argv[3]), ticks = std::atoi(argv[4]);
        TrajoptSolver solver(model, goal, N, 0.01);
        // the simulator's home configuration (src/sim/sim/sim_node.cpp:196), at rest
        std::vector<double> state = {1.57994932, 0.06313131, -1.18071498, 1.09272369, -0.6255293, -0.01895007, 0, 0, 0, 0, 0, 0};
        std::printf("%s\n", i7_version());
        for (int k = 0; k < ticks && !solver.isTrajectoryComplete(); ++k) {
            const double now = 0.005 + 0.013 * k;              // state messages every 13 ms: one or two knots per tick
            solver.shiftTrajectory(state, now);
            const std::string msg = solver.runTrajoptIteration();
            const std::vector<double>& xu = solver.getOptimizedTrajectory();
            std::printf("tick %d offset %u | %s |", k, solver.getTrajectoryOffset(), msg.c_str());
            for (int i = 12; i < 30; ++i) std::printf(" %.17g", xu[i]);       // u_0 and x_1
            std::printf("\n");
            state.assign(xu.begin() + 18, xu.begin() + 30);    // the plant follows the plan (a timing / plumbing run)
        }
    } catch (const std::exception& e) {
        std::fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
    return 0;
}
