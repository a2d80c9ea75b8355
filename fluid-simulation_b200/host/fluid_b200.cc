// fluid_b200.cc -- host driver that replaces main() of the reference's fluid.cc (:1151-1514): same
// scene, same step loop, same stdout lines, same output files; the loop body (:1378-1490) is one
// call into libflipb200.so (include/flipb200.h).  Built and run by ./run.sh fluid.
//
// Scene (all literals of fluid.cc): index box [-60,60]^3 (:1159), solid shell |c| > 58 (:1264),
// fluid cube [-20,20]^3 (:1176), UniformPointScatter with 10 points per voxel volume and
// mt19937(0) (:1348-1350), PointList::add's |p| < bound-2 filter (:841), gravity (0,-10,0), dx = 1,
// dt0 = 0.1, 500 steps (:1357-1368).  Environment overrides (not in the reference, which has no
// flags): FLUID_STEPS, FLUID_VDB=0|1, FLUID_VDB_ZIP=0|1, FLUID_STEP_FILES=0|1, FLUID_PRECOND=0|1, and
// the scenes the reference keeps commented out: FLUID_SCENE=normal|pea|side|stable (:1176-1216),
// FLUID_OBSTACLE=two_blocks|three_blocks|big_wall (:1303-1345), FLUID_RESEED=k (the re-seeding hook
// of :1374-1382 / :1495-1497: every k-th step new particles enter the fluid boxes with the PIC
// velocity of interpFromGrid, flipb200_emit_box).
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
struct IBox { int lo[3], hi[3]; };   // inclusive, reference coordinates
static void scatter(const IBox& b, float ppv, unsigned seed, int bound, std::vector<double> p[3]) {
  const uint64_t nx = (uint64_t)(b.hi[0] - b.lo[0] + 1), ny = (uint64_t)(b.hi[1] - b.lo[1] + 1), nz = (uint64_t)(b.hi[2] - b.lo[2] + 1);
  const uint64_t nvox = nx * ny * nz;
  const uint64_t target = (uint64_t)(ppv * 1.0 * 1.0 * 1.0) * nvox;          // PointScatter.h:150
  std::mt19937 gen(seed);
  std::uniform_int_distribution<uint64_t> pick(0, nvox - 1);                 // math::RandInt, Math.h:174
  std::vector<uint64_t> ids(target);
  for (uint64_t i = 0; i < target; ++i) ids[i] = pick(gen);
  std::sort(ids.begin(), ids.end());                                          // tbb::parallel_sort, :161
  std::uniform_real_distribution<double> r01(0.0, 1.0);                      // math::Rand01, Math.h:139
  for (uint64_t i = 0; i < target; ++i) {
    const uint64_t id = ids[i];
    const double c[3] = {(double)(b.lo[0] + (int64_t)(id / (ny * nz))), (double)(b.lo[1] + (int64_t)((id / nz) % ny)),
                         (double)(b.lo[2] + (int64_t)(id % nz))};
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
  // fluid boxes (fluid.cc:1176 and the commented alternatives :1178-1216)
  const std::string scene = getenv("FLUID_SCENE") ? getenv("FLUID_SCENE") : "normal";
  std::vector<IBox> fluid;
  if (scene == "pea") { fluid.push_back({{-1, -1, -1}, {1, 1, 1}}); fluid.push_back({{-1, 6, -1}, {1, 9, 1}}); }
  else if (scene == "side") fluid.push_back({{-57, -57, -57}, {57, -40, -40}});
  else if (scene == "stable") fluid.push_back({{-57, -57, -57}, {57, -55, 57}});
  else fluid.push_back({{-20, -20, -20}, {20, 20, 20}});
  // interior obstacles (:1303-1329 inside the shell loop, :1334-1345 after it)
  const std::string obst = getenv("FLUID_OBSTACLE") ? getenv("FLUID_OBSTACLE") : "";
  std::vector<IBox> blocks;
  if (obst == "two_blocks") { blocks.push_back({{-11, -58, -3}, {-6, -8, 3}}); blocks.push_back({{6, -58, -3}, {11, -8, 3}}); }
  else if (obst == "three_blocks") { blocks.push_back({{-11, -58, -3}, {-7, -8, 3}}); blocks.push_back({{-2, -58, -3}, {2, -8, 3}}); blocks.push_back({{7, -58, -3}, {11, -8, 3}}); }
  else if (obst == "big_wall") blocks.push_back({{-58, -58, -30}, {58, -50, -25}});
  std::vector<uint8_t> solid;
  if (!blocks.empty()) {
    solid.assign((size_t)n * n * n, 0);
    for (const IBox& b : blocks)
      for (int z = b.lo[2]; z <= b.hi[2]; ++z) for (int y = b.lo[1]; y <= b.hi[1]; ++y) for (int x = b.lo[0]; x <= b.hi[0]; ++x)
        solid[((size_t)(z + bound) * n + (y + bound)) * n + (x + bound)] = 1;
  }
  const int reseed = env_int("FLUID_RESEED", 0);

  flipb200_ctx* ctx = nullptr;
  int rc = flipb200_create(&prm, solid.empty() ? nullptr : solid.data(), &ctx);   // shell |c| > 58 is built by the library (:1256-1331)
  if (rc) { fprintf(stderr, "flipb200_create: %s\n", flipb200_strerror(rc)); return 1; }

  std::vector<double> p[3], v[3];
  for (size_t b = 0; b < fluid.size(); ++b) scatter(fluid[b], 10.f, (unsigned)b, bound, p);   // fluid.cc:1176,1348-1350
  for (int a = 0; a < 3; ++a) v[a].assign(p[a].size(), 0.0);                // velocities start at 0 (:841)
  rc = flipb200_set_particles(ctx, p[0].data(), p[1].data(), p[2].data(), v[0].data(), v[1].data(), v[2].data(),
                              (int64_t)p[0].size());
  if (rc) { fprintf(stderr, "flipb200_set_particles: %s\n", flipb200_strerror(rc)); return 1; }

  GridSink* sink = vdb ? new GridSink("mygrids.vdb", steps, zip, n) : nullptr;   // fluid.cc:1364-1366,1508
  double dt = 0.1, simulationTime = 0;                                      // fluid.cc:1363,1367
  const size_t N = (size_t)n * n * n;
  for (int i = 0; i < steps; ++i) {
    if (reseed > 0 && i > 0 && i % reseed == 0)                             // :1379-1382 / :1495-1497 (commented out there)
      for (const IBox& b : fluid) {
        const int32_t lo[3] = {b.lo[0] + bound, b.lo[1] + bound, b.lo[2] + bound}, hi[3] = {b.hi[0] + bound, b.hi[1] + bound, b.hi[2] + bound};
        rc = flipb200_emit_box(ctx, lo, hi, (uint64_t)(i + 1));             // mtRand(i + 1), :1375
        if (rc) { fprintf(stderr, "flipb200_emit_box: %s\n", flipb200_strerror(rc)); return 1; }
      }
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
