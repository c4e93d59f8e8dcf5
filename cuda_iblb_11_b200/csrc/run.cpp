// Driver-level run: what the reference's main() does around the time-step loop — the Data/ tree,
// SimLog.txt, the flux file, and (BigData) the fluid/cilia snapshots.  Replaces main.cu:589-631,
// 758-790, 936-1005, 1030-1060.  Pure host C++ on top of the C ABI; the numbers go through
// default-formatted ostreams exactly as in the reference, so the text matches digit for digit.
#include <sys/stat.h>
#include <sys/types.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <functional>
#include <memory>
#include <mutex>
#include <thread>
#include <ctime>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>

#include "iblb.h"

namespace iblb {
void set_error(const std::string &msg);
}

namespace {

// main.cu:255-261
template <typename T>
std::string to_string_3(const T v, const int n = 3) {
    std::ostringstream out;
    out << std::setprecision(n) << v;
    return out.str();
}

void mkdir_p(const std::string &path) {
    std::string cur;
    for (size_t i = 0; i < path.size(); i++) {
        cur += path[i];
        if (path[i] == '/' || i + 1 == path.size()) mkdir(cur.c_str(), 0777);
    }
}

double now_seconds() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch())
        .count();
}

// one fluid + cilia snapshot on its way to the text files
struct Snapshot {
    std::vector<double> rho, u;
    std::vector<float> sm, us;
    std::vector<int> eps;
};

// Writer thread: the reference formats and flushes one line per node on the
// thread that drives the GPU (main.cu:954-994, `endl` per line), so BigData=1 stalls its time loop
// for the whole write.  Here the loop hands the snapshot over and steps on; at most two snapshots
// are in flight (submit blocks beyond that), files appear in iteration order.
class SnapshotWriter {
  public:
    SnapshotWriter() : ok_(true), stop_(false), worker_([this] { loop(); }) {}
    ~SnapshotWriter() { finish(); }
    void submit(std::function<bool()> job) {
        std::unique_lock<std::mutex> lk(m_);
        cv_.wait(lk, [this] { return jobs_.size() < 2; });
        jobs_.push_back(std::move(job));
        cv_.notify_all();
    }
    // drain and join; false when any file could not be written
    bool finish() {
        {
            std::unique_lock<std::mutex> lk(m_);
            stop_ = true;
            cv_.notify_all();
        }
        if (worker_.joinable()) worker_.join();
        return ok_;
    }

  private:
    void loop() {
        for (;;) {
            std::function<bool()> job;
            {
                std::unique_lock<std::mutex> lk(m_);
                cv_.wait(lk, [this] { return stop_ || !jobs_.empty(); });
                if (jobs_.empty()) return;
                job = jobs_.front();
            }
            const bool ok = job();
            std::unique_lock<std::mutex> lk(m_);
            if (!ok) ok_ = false;
            jobs_.pop_front();
            cv_.notify_all();
        }
    }
    std::mutex m_;
    std::condition_variable cv_;
    std::deque<std::function<bool()>> jobs_;
    bool ok_, stop_;
    std::thread worker_;
};

}  // namespace

namespace {

// Columns [x0, x0 + w) of slab r of n: equal split of the channel in units of `align` columns
// (c_space with cilia, so that a cilium's base never sits on a seam), remainder to the low ranks.
bool slab_range(int r, int n, unsigned XDIM, unsigned align, unsigned &x0, unsigned &w) {
    if (align == 0) align = 1;
    const unsigned units = XDIM / align;
    if (units * align != XDIM || units < (unsigned)n) return false;
    const unsigned base = units / n, rem = units % n;
    w = (base + ((unsigned)r < rem ? 1 : 0)) * align;
    x0 = ((unsigned)r * base + std::min((unsigned)r, rem)) * align;
    return true;
}

// The channel of one run: the whole of it on one GPU, or x-slabs on several GPUs of this process
// (iblb_connect_local: peer stores into the neighbours' ghost columns, CUDA events between the
// slabs' streams), advanced in lockstep by this one host thread.  The reference's analogue is its
// device selection (main.cu:395-396): it is single-GPU.
struct Channel {
    std::vector<iblb_solver *> slab;
    std::vector<unsigned> x0, w;
    unsigned XDIM = 0, YDIM = 0;

    ~Channel() {
        for (iblb_solver *s : slab) iblb_destroy(s);
    }
    int create(const iblb_params &p, int n_gpus, const char *restart_path) {
        XDIM = p.XDIM;
        YDIM = p.YDIM;
        if (n_gpus <= 1) {
            iblb_solver *s = nullptr;
            int rc = iblb_create(&p, &s);
            if (rc) return rc;
            slab.push_back(s);
            x0.push_back(0);
            w.push_back(p.XDIM);
            return restart_path ? iblb_load_checkpoint(s, restart_path) : IBLB_OK;
        }
        for (int r = 0; r < n_gpus; r++) {
            iblb_params q = p;
            unsigned a = 0, b = 0;
            if (!slab_range(r, n_gpus, p.XDIM, p.c_num ? p.c_space : 1, a, b)) {
                iblb::set_error("iblb_run: the channel cannot be cut into " + std::to_string(n_gpus) +
                                " slabs of whole cilium spacings");
                return IBLB_EINVAL;
            }
            q.slab_x0 = a;
            q.slab_w = b;
            // slab r on GPU (first + r); with fewer GPUs than slabs they share (same results: the
            // slabs are ordered by events, not by where they run)
            const int ndev = iblb_device_count();
            q.device = ((p.device > 0 ? p.device : 0) + r) % (ndev > 0 ? ndev : 1);
            iblb_solver *s = nullptr;
            int rc = iblb_create(&q, &s);
            if (rc) return rc;
            slab.push_back(s);
            x0.push_back(a);
            w.push_back(b);
            if (restart_path) {  // one file per slab, before the slabs are connected
                rc = iblb_load_checkpoint(s, (std::string(restart_path) + "." + std::to_string(r)).c_str());
                if (rc) return rc;
            }
        }
        if (restart_path) {
            // One file per slab: each file is complete or absent, but a run killed while save() was
            // renaming can leave files of two different iterations side by side.  Slabs of
            // different iterations must never be stepped together.
            for (int r = 1; r < n_gpus; r++)
                if (iblb_iteration(slab[r]) != iblb_iteration(slab[0])) {
                    iblb::set_error("iblb_run: the checkpoint files " + std::string(restart_path) +
                                    ".0 .. ." + std::to_string(n_gpus - 1) +
                                    " are not one set: slab 0 is at iteration " +
                                    std::to_string(iblb_iteration(slab[0])) + ", slab " +
                                    std::to_string(r) + " at " + std::to_string(iblb_iteration(slab[r])) +
                                    " (the run was stopped while it was saving; use an older set)");
                    return IBLB_EINVAL;
                }
        }
        for (int r = 0; r < n_gpus; r++) {
            int rc = iblb_connect_local(slab[r], slab[(r + n_gpus - 1) % n_gpus], slab[(r + 1) % n_gpus]);
            if (rc) return rc;
        }
        return IBLB_OK;
    }
    unsigned iteration() const { return iblb_iteration(slab[0]); }
    int step(unsigned n) {
        if (slab.size() == 1) return iblb_step(slab[0], n);
        for (unsigned k = 0; k < n; k++)  // lockstep: every slab one iteration, then the next
            for (iblb_solver *s : slab) {
                int rc = iblb_step(s, 1);
                if (rc) return rc;
            }
        return IBLB_OK;
    }
    int sync() {
        for (iblb_solver *s : slab) {
            int rc = iblb_sync(s);
            if (rc) return rc;
        }
        return IBLB_OK;
    }
    // rho[size], u[2*size] of the whole channel in the reference's layout (main.cu:944-952)
    int fields(double *rho, double *u) {
        if (slab.size() == 1) return iblb_get_fields(slab[0], rho, u);
        const size_t size = (size_t)XDIM * YDIM;
        std::vector<double> r, v;
        for (size_t i = 0; i < slab.size(); i++) {
            const size_t ls = (size_t)w[i] * YDIM;
            r.resize(ls);
            v.resize(2 * ls);
            int rc = iblb_get_fields(slab[i], r.data(), v.data());
            if (rc) return rc;
            for (unsigned y = 0; y < YDIM; y++) {
                const size_t g = (size_t)y * XDIM + x0[i], l = (size_t)y * w[i];
                memcpy(rho + g, r.data() + l, sizeof(double) * w[i]);
                memcpy(u + g, v.data() + l, sizeof(double) * w[i]);
                memcpy(u + size + g, v.data() + ls + l, sizeof(double) * w[i]);
            }
        }
        return IBLB_OK;
    }
    // every slab evaluates all cilia (prescribed kinematics): slab 0 speaks for them
    int markers(float *pos, float *vel, int *eps) {
        return iblb_get_markers(slab[0], pos, vel, nullptr, eps);
    }
    // the flux plane lives in exactly one slab; the others hold 0
    int flux(double *Q) {
        double q = 0.;
        for (iblb_solver *s : slab) {
            double v = 0.;
            int rc = iblb_get_flux(s, &v);
            if (rc) return rc;
            q += v;
        }
        *Q = q;
        return IBLB_OK;
    }
    int save(const char *path) {
        if (slab.size() == 1) return iblb_save_checkpoint(slab[0], path);
        // Two phases, so that the window in which the set is mixed is N renames and not N multi-GB
        // writes: every slab first goes to <path>.<i>.new (iblb_save_checkpoint itself writes to a
        // temporary and renames), and only when all of them are on disk do they replace the old
        // set.  What is left of that window is caught on load (create() compares the iterations).
        for (size_t i = 0; i < slab.size(); i++) {
            int rc = iblb_save_checkpoint(slab[i], (std::string(path) + "." + std::to_string(i) + ".new").c_str());
            if (rc) return rc;
        }
        for (size_t i = 0; i < slab.size(); i++) {
            const std::string to = std::string(path) + "." + std::to_string(i), from = to + ".new";
            if (rename(from.c_str(), to.c_str()) != 0) {
                iblb::set_error("iblb_run: cannot rename " + from + " to " + to);
                return IBLB_EIO;
            }
        }
        return IBLB_OK;
    }
};

}  // namespace

extern "C" int iblb_slab_range(int rank, int n_slabs, unsigned XDIM, unsigned align, unsigned *x0,
                               unsigned *w) {
    if (!x0 || !w || n_slabs < 1 || rank < 0 || rank >= n_slabs ||
        !slab_range(rank, n_slabs, XDIM, align, *x0, *w)) {
        iblb::set_error("iblb_slab_range: XDIM cannot be cut into that many slabs of whole `align` columns");
        return IBLB_EINVAL;
    }
    return IBLB_OK;
}

extern "C" int iblb_run_gpus(const iblb_params *pp, const char *root, int n_gpus) {
    using namespace std;
    if (!pp) return IBLB_EINVAL;
    iblb_params p = *pp;
    const unsigned XDIM = p.XDIM, YDIM = p.YDIM, c_num = p.c_num, c_fraction = p.c_fraction;
    const unsigned ITERATIONS = p.ITERATIONS, INTERVAL = p.INTERVAL;
    const size_t size = (size_t)XDIM * YDIM;
    const int mpc = p.LENGTH ? (int)p.LENGTH : 96;  // markers per cilium (main.cu:83,442)
    const int Ns = mpc * (int)c_num;

    // ---- directories, main.cu:589-613 (root is configurable; the reference hard-codes it) ----
    string output_data = root ? root : "Data/";
    if (!output_data.empty() && output_data.back() != '/') output_data += '/';
    string raw_data = output_data + "Raw/" + to_string(c_num) + "/" + to_string(c_fraction) + "/";
    string cilia_data = output_data + "Cilia/" + to_string(c_num) + "/" + to_string(c_fraction) + "/";
    mkdir_p(raw_data);
    mkdir_p(cilia_data);
    mkdir_p(output_data + "Flux/");
    string flux = output_data + "/Flux/" + to_string(c_fraction) + "_" + to_string(c_num) + "_" +
                  to_string(p.c_space) + "_" + to_string_3(p.Re) + "_" + to_string_3(p.T_num) +
                  "x" + to_string_3(p.T_pow) + "-flux.dat";
    string parameters = raw_data + "/SimLog.txt";

    // checkpoint / restart (extension; environment only, the argv surface stays the reference's):
    //   IBLB_CHECKPOINT=<file>  written after every output interval and at the end of the run
    //   IBLB_RESTART=<file>     continue from it: the flux file is appended to, not truncated
    const char *ckpt_path = getenv("IBLB_CHECKPOINT");
    const char *restart_path = getenv("IBLB_RESTART");
    if (ckpt_path && !*ckpt_path) ckpt_path = nullptr;
    if (restart_path && !*restart_path) restart_path = nullptr;

    ofstream fsB, fsC;
    fsB.open(flux.c_str(), restart_path ? ofstream::app : ofstream::trunc);
    if (!fsB) {
        iblb::set_error("cannot write " + flux);
        return IBLB_EIO;
    }
    fsB.close();

    time_t rawtime;
    time(&rawtime);
    struct tm *timeinfo = localtime(&rawtime);
    cout << asctime(timeinfo) << endl;
    cout << "Initialising...\n";

    Channel ch;  // destroys its solvers on every way out
    int rc = ch.create(p, n_gpus, restart_path);
    if (rc) return rc;

    // ---- SimLog header, main.cu:761-790 ----
    const double l_0 = 0.000006, t_0 = 0.067;
    const double dx = 1. / p.LENGTH, dt = 1. / p.T;
    fsC.open(parameters.c_str(), ofstream::trunc);
    fsC << asctime(timeinfo) << endl;
    fsC << "Size: " << XDIM << "x" << YDIM << endl;
    fsC << "Iterations: " << ITERATIONS << endl;
    fsC << "Reynolds Number: " << p.Re << endl;
    fsC << "Relaxation times: " << p.TAU << ", " << p.TAU2 << endl;
    fsC << "Spatial step: " << dx * l_0 << "m" << endl;
    fsC << "Time step: " << dt * t_0 << "s" << endl;
    fsC << "Mach number: " << p.Ma << endl;
    fsC << "Phase Step: " << c_fraction << "/" << c_num << endl;
    if (p.BigData) fsC << "\nBig Data is ON" << endl;
    else fsC << "\nBig Data is OFF" << endl;
    if (p.ShARC) fsC << "Running on ShARC" << endl;
    else fsC << "Running on local GPU" << endl;

    cout << "Running Simulation...\n";
    const double start = now_seconds();

    SnapshotWriter writer;  // BigData text files are written behind the step loop
    long long nonfinite_it = -1;  // first output iteration at which Q was not finite

    // iteration `it` (0-based) is complete after it+1 steps; outputs follow iterations with
    // it % INTERVAL == 0 (main.cu:938); everything in between runs without host involvement
    unsigned done = ch.iteration();  // 0, or where the checkpoint was taken
    while (done < ITERATIONS && rc == IBLB_OK) {
        unsigned it_out;  // next iteration index whose completion triggers output
        if (INTERVAL == 0) it_out = ITERATIONS;  // no periodic output
        else it_out = ((done + INTERVAL - 1) / INTERVAL) * INTERVAL;
        unsigned target = (it_out < ITERATIONS) ? it_out + 1 : ITERATIONS;
        rc = ch.step(target - done);
        if (rc) break;
        done = target;
        const unsigned it = done - 1;
        if (INTERVAL != 0 && it % INTERVAL == 0) {
            if (p.BigData) {
                // the state is read back here; formatting and writing the text files (the slow
                // part: one line per node) happen on the writer thread while the loop steps on
                auto snap = make_shared<Snapshot>();
                snap->rho.resize(size);
                snap->u.resize(2 * size);
                snap->sm.resize(2 * (size_t)Ns + 2);
                snap->us.resize(2 * (size_t)Ns + 2);
                snap->eps.resize((size_t)Ns + 1);
                rc = ch.fields(snap->rho.data(), snap->u.data());
                if (rc) break;
                rc = ch.markers(snap->sm.data(), snap->us.data(), snap->eps.data());
                if (rc) break;
                const string fluid_file = raw_data + to_string(it) + "-fluid.dat";
                const string cilia_file = cilia_data + to_string(it) + "-cilia.dat";
                const double x_scale = p.x_scale, s_scale = p.s_scale;
                writer.submit([=]() -> bool {
                    const vector<double> &rho = snap->rho, &u = snap->u;
                    const vector<float> &sm = snap->sm, &us = snap->us;
                    const vector<int> &eps = snap->eps;
                    ofstream f(fluid_file.c_str());
                    for (size_t j = 0; j < size; j++) {  // main.cu:958-969
                        const int x = (int)(j % XDIM);
                        const int y = (int)((j - j % XDIM) / XDIM);
                        const double ab = sqrt(u[j] * u[j] + u[size + j] * u[size + j]);
                        f << x * x_scale << "\t" << y * x_scale << "\t" << u[j] * s_scale << "\t"
                          << u[size + j] * s_scale << "\t" << ab * s_scale << "\t" << rho[j] << "\n";
                        if (x == (int)XDIM - 1) f << "\n";
                    }
                    f.close();
                    bool ok = !f.fail();
                    f.open(cilia_file.c_str());
                    for (int k = 0; k < Ns; k++) {  // main.cu:986-992
                        f << sm[2 * k + 0] * x_scale << "\t" << sm[2 * k + 1] * x_scale << "\t"
                          << us[2 * k + 0] * s_scale << "\t" << us[2 * k + 1] * s_scale << "\t"
                          << eps[k] << "\n";
                        if (k % mpc == mpc - 1 || sm[2 * k + 0] > XDIM - 1 || sm[2 * k + 0] < 1) f << "\n";
                    }
                    f.close();
                    return ok && !f.fail();
                });
            }
            double Q = 0.;
            rc = ch.flux(&Q);
            if (rc) break;
            fsB.open(flux.c_str(), ofstream::app);
            fsB << it * p.t_scale << "\t" << Q * p.x_scale << endl;  // main.cu:1002
            fsB.close();
            // The reference's coupling is unstable for phase-lagged cilia (its own build overflows
            // near iteration 42 on cilia6.sh:7) and it writes NaN files to the
            // end without a word; so does this driver, but it says so.
            if (!std::isfinite(Q) && nonfinite_it < 0) {
                nonfinite_it = (long long)it;
                cerr << "app: warning: the flux is non-finite at iteration " << it
                     << " (checked every " << INTERVAL << "); the run continues as the reference's does"
                     << endl;
            }
        }
        if (ckpt_path && (rc = ch.save(ckpt_path)) != IBLB_OK) break;
        if (it == INTERVAL && INTERVAL != 0) {  // main.cu:1007-1022
            const double cycle = now_seconds();
            time_t p_runtime = (time_t)((cycle - start) * (ITERATIONS / INTERVAL));
            time_t p_end = rawtime + p_runtime;
            timeinfo = localtime(&p_end);
            cout << "\nCompletion time: " << asctime(timeinfo) << endl;
            fsC << "\nCompletion time: " << asctime(timeinfo) << endl;
            fsC.close();
        }
    }
    if (!writer.finish() && rc == IBLB_OK) {
        iblb::set_error("cannot write the snapshot files under " + output_data);
        rc = IBLB_EIO;
    }
    if (rc == IBLB_OK) {
        double Q = 0.;
        rc = ch.flux(&Q);
        fsB.open(flux.c_str(), ofstream::app);
        fsB << ITERATIONS * p.t_scale << "\t" << Q * p.x_scale << endl;  // main.cu:1030-1034
        fsB.close();
    }
    string keep = iblb_last_error();
    ch.sync();
    const double runtime = now_seconds() - start;  // main.cu:1038-1057
    int hours = 0, mins = 0;
    if (runtime > 3600) hours = (int)nearbyint(runtime / 3600 - 0.5);
    if (runtime > 60) mins = (int)nearbyint((runtime - hours * 3600) / 60 - 0.5);
    const double secs = runtime - hours * 3600 - mins * 60;
    if (fsC.is_open()) fsC.close();
    fsC.open(parameters.c_str(), ofstream::app);
    fsC << "Total runtime: ";
    if (hours < 10) fsC << 0;
    fsC << hours << ":";
    if (mins < 10) fsC << 0;
    fsC << mins << ":";
    if (secs < 10) fsC << 0;
    fsC << secs << endl;
    // additions of this implementation (after the reference's last line)
    fsC << "MLUPS: " << (runtime > 0 ? (double)size * done / runtime / 1e6 : 0.) << endl;
    if (nonfinite_it >= 0) fsC << "Non-finite flux from output iteration: " << nonfinite_it << endl;
    if (ch.slab.size() > 1) fsC << "GPUs: " << ch.slab.size() << " x-slabs" << endl;
    fsC.close();
    if (rc) iblb::set_error(keep);
    return rc;
}

// main() as a whole on one GPU, or on IBLB_GPUS GPUs of this process when that is set
extern "C" int iblb_run(const iblb_params *p, const char *root) {
    const char *e = getenv("IBLB_GPUS");
    return iblb_run_gpus(p, root, e ? atoi(e) : 1);
}
