// fluid_b200.cc -- host driver that replaces main() of the reference's fluid.cc (:1151-1514): same
// scene, same step loop, same stdout lines, same output files; the loop body (:1378-1490) is one
// call into libflipb200.so (include/flipb200.h).  Built and run by ./run.sh fluid.
//
// Scene (all literals of fluid.cc): index box [-60,60]^3 (:1159), solid shell |c| > 58 (:1264),
// fluid cube [-20,20]^3 (:1176), UniformPointScatter with 10 points per voxel volume and
// mt19937(0) (:1348-1350), PointList::add's |p| < bound-2 filter (:841), gravity (0,-10,0), dx = 1,
// dt0 = 0.1, 500 steps (:1357-1368).  Environment overrides (not in the reference, which has no
// flags): FLUID_STEPS, FLUID_VDB=0|1, FLUID_VDB_ZIP=0|1, FLUID_STEP_FILES=0|1, FLUID_PRECOND=0|1.
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <ctime>
#include <cstring>
#include <iostream>
#include <random>
#include <string>
#include <vector>
#include <algorithm>
#include <thread>
#include <mutex>
#include <condition_variable>
#include <deque>
#include <sys/stat.h>
#include "flipb200.h"
#include "vdb_writer.h"

static int env_int(const char* k, int d) { const char* v = getenv(k); return v ? atoi(v) : d; }

// UniformPointScatter restated (openvdb/tools/PointScatter.h:143-185, addPoint :420-428): target
// = Index64(ppv * voxel volume) * voxelCount sorted uniform voxel picks, one Rand01 offset per axis
// in [c-0.5, c+0.5).  Voxel visiting order is x-major here (the VDB tree's ValueOnCIter order and
// libstdc++'s distributions make the reference's exact sequence toolchain-dependent).
static void scatter(int lo, int hi, float ppv, unsigned seed, int bound, std::vector<double> p[3]) {
  const uint64_t n1 = (uint64_t)(hi - lo + 1), nvox = n1 * n1 * n1;
  const uint64_t target = (uint64_t)(ppv * 1.0 * 1.0 * 1.0) * nvox;          // PointScatter.h:150
  std::mt19937 gen(seed);
  std::uniform_int_distribution<uint64_t> pick(0, nvox - 1);                 // math::RandInt, Math.h:174
  std::vector<uint64_t> ids(target);
  for (uint64_t i = 0; i < target; ++i) ids[i] = pick(gen);
  std::sort(ids.begin(), ids.end());                                          // tbb::parallel_sort, :161
  std::uniform_real_distribution<double> r01(0.0, 1.0);                      // math::Rand01, Math.h:139
  for (uint64_t i = 0; i < target; ++i) {
    const uint64_t id = ids[i];
    const double c[3] = {(double)(lo + (int64_t)(id / (n1 * n1))), (double)(lo + (int64_t)((id / n1) % n1)),
                         (double)(lo + (int64_t)(id % n1))};
    double q[3];
    for (int a = 0; a < 3; ++a) q[a] = c[a] - 0.5 + r01(gen);                // coord - 0.5 + getRand(), :420-428
    if (std::fabs(q[0]) < bound - 2 && std::fabs(q[1]) < bound - 2 && std::fabs(q[2]) < bound - 2)   // PointList::add, fluid.cc:841
      for (int a = 0; a < 3; ++a) p[a].push_back(q[a]);
  }
}

// background writer: compressing 4096 leaves per grid with zlib must not stall the step loop
class GridSink {
 public:
  GridSink(const std::string& path, int count, bool zip, int n) : w_(path, count, zip), n_(n), stop_(false) {
    th_ = std::thread([this] { run(); });
  }
  void push(std::vector<float>&& g) {
    std::unique_lock<std::mutex> l(m_);
    cv_.wait(l, [this] { return q_.size() < 4; });
    q_.push_back(std::move(g));
    cv_.notify_all();
  }
  bool finish() {
    { std::unique_lock<std::mutex> l(m_); stop_ = true; cv_.notify_all(); }
    th_.join();
    return w_.close();
  }
  bool ok() const { return w_.ok(); }
 private:
  void run() {
    for (;;) {
      std::vector<float> g;
      {
        std::unique_lock<std::mutex> l(m_);
        cv_.wait(l, [this] { return stop_ || !q_.empty(); });
        if (q_.empty()) return;
        g = std::move(q_.front()); q_.pop_front();
        cv_.notify_all();
      }
      flipvdb::DenseGrid d; d.lo[0] = d.lo[1] = d.lo[2] = -(n_ / 2); d.dim[0] = d.dim[1] = d.dim[2] = n_;
      d.data = g.data(); d.name = "";
      w_.write_grid(d);
    }
  }
  flipvdb::Writer w_; int n_; bool stop_;
  std::thread th_; std::mutex m_; std::condition_variable cv_; std::deque<std::vector<float>> q_;
};

int main(int, char**) {
  const clock_t start = clock();                                            // fluid.cc:20
  const int bound = 60, n = 2 * bound + 1;                                  // box [-60,60]
  const int steps = env_int("FLUID_STEPS", 500);                            // fluid.cc:1368
  const bool vdb = env_int("FLUID_VDB", 1) != 0, zip = env_int("FLUID_VDB_ZIP", 1) != 0;
  struct stat sb;
  const bool step_files = env_int("FLUID_STEP_FILES", 1) != 0 && stat("simulation", &sb) == 0 && S_ISDIR(sb.st_mode);

  flipb200_params prm;
  flipb200_default_params(&prm);                                            // the literals of fluid.cc
  prm.precond = env_int("FLUID_PRECOND", prm.precond);
  flipb200_ctx* ctx = nullptr;
  int rc = flipb200_create(&prm, nullptr, &ctx);                            // shell |c| > 58 is built by the library (:1256-1331)
  if (rc) { fprintf(stderr, "flipb200_create: %s\n", flipb200_strerror(rc)); return 1; }

  std::vector<double> p[3], v[3];
  scatter(-20, 20, 10.f, 0u, bound, p);                                     // fluid.cc:1176,1348-1350
  for (int a = 0; a < 3; ++a) v[a].assign(p[a].size(), 0.0);                // velocities start at 0 (:841)
  rc = flipb200_set_particles(ctx, p[0].data(), p[1].data(), p[2].data(), v[0].data(), v[1].data(), v[2].data(),
                              (int64_t)p[0].size());
  if (rc) { fprintf(stderr, "flipb200_set_particles: %s\n", flipb200_strerror(rc)); return 1; }

  GridSink* sink = vdb ? new GridSink("mygrids.vdb", steps, zip, n) : nullptr;   // fluid.cc:1364-1366,1508
  double dt = 0.1, simulationTime = 0;                                      // fluid.cc:1363,1367
  const size_t N = (size_t)n * n * n;
  for (int i = 0; i < steps; ++i) {
    std::cout << "2" << std::endl;                                          // fluid.cc:1383
    std::cout << "3" << std::endl;                                          // :1385
    std::cout << "DT " << dt << std::endl;                                  // :1386
    std::cout << "Before" << std::endl;                                     // :1456
    flipb200_stats st;
    rc = flipb200_step(ctx, &dt, &st);                                      // body :1378-1490
    if (rc) { fprintf(stderr, "flipb200_step: %s\n", flipb200_strerror(rc)); return 1; }
    std::cout << "After" << std::endl;                                      // :1486
    std::cout << "DT " << dt << std::endl;                                  // :1491
    std::cout << "Error:\t" << st.last_error << std::endl;                  // :1499
    std::cout << "Iteration:\t" << i + 1 << std::endl;                      // :1500
    simulationTime += dt;                                                   // :1501
    std::cout << "Time delta:\t" << simulationTime << std::endl;            // :1502
    if (vdb || step_files) {
      std::vector<float> out(N);                                            // outputGrid (:1434-1451)
      rc = flipb200_get_grid(ctx, FLIPB200_GRID_WEIGHT, out.data(), N * sizeof(float));
      if (rc) { fprintf(stderr, "flipb200_get_grid: %s\n", flipb200_strerror(rc)); return 1; }
      if (step_files) {                                                     // simulation/mygrids<i>.vdb (:1371,1503)
        flipvdb::Writer w("simulation/mygrids" + std::to_string(i) + ".vdb", 1, zip);
        flipvdb::DenseGrid d; d.lo[0] = d.lo[1] = d.lo[2] = -bound; d.dim[0] = d.dim[1] = d.dim[2] = n; d.data = out.data();
        w.write_grid(d); w.close();
      }
      if (sink) sink->push(std::move(out));
    }
  }
  bool ok = true;
  if (sink) { ok = sink->finish(); delete sink; }                           // file.write(grids), :1508-1509
  flipb200_destroy(ctx);
  const double cpu_time_used = ((double)(clock() - start)) / CLOCKS_PER_SEC;   // :1511-1512
  std::cout << "Time Taken " << cpu_time_used / 60 << " minutes" << std::endl;   // :1513
  return ok ? 0 : 2;
}
