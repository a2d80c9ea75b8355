/*
 * ref_bridge.cc -- TEST INFRASTRUCTURE (see oracle/flip_oracle.h): the REFERENCE ITSELF behind a C ABI.
 *
 * This translation unit textually includes /root/reference/fluid.cc WHERE IT LIES (never copied; found
 * through -I$(REF)), unmodified, with `main` renamed to `ref_main`, on top of the dense-array OpenVDB/TBB
 * stand-in in oracle/shim/ and the reference's own vendored Eigen 3.3.4.  Everything the tests compare
 * the restatement (oracle/flip_oracle.cc) against is therefore computed by the reference's own functions:
 *
 *   function level  spline :22-37, PointList::P2Gtransfer :1106-1148 (p2gCatmullRom :265-299),
 *                   PointList::interpolate :843-882, setRHS :414-479, setDiver :566-610, setA :304-412,
 *                   setA2 :481-541 (+ the reference's cg.compute / cg.solve, :1473-1474), setOnlyB :543-564,
 *                   velUpdate :612-703, PointList::FLIPadvect :972-1038 (CatmullRomFLIP :210-263,
 *                   getVelocity :59-70), extrapolate :705-802, PointList::interpFromGrid :883-895
 *                   (clampedCatmullRom :125-207), PointList::advect :904-971
 *   whole program   ref_main() = fluid.cc's main() :1151-1514 run for K steps of its own 500-step loop on its
 *                   own scene (the shim's io::File::write hook snapshots the state at the end of step K and
 *                   leaves the loop by an exception)
 *
 * Built by `make -C oracle fluidref` into oracle/_ref/libfluid_ref.so (git-ignored, travels prebuilt).
 * Grid arrays cross this ABI as dense 121^3 arrays, idx = ((z+60)*121 + (y+60))*121 + (x+60) -- the
 * layout of oracle/flip_oracle.h; the box [-60,60]^3 and the 58 shell are literals of fluid.cc.
 */
#define main ref_main
#include <fluid.cc>
#undef main

#include <cstring>
#include <sstream>

namespace {

const int RB = 60, RD = 121;
const size_t RN = (size_t)RD * RD * RD;
inline size_t ridx(int x, int y, int z) { return ((size_t)(z + RB) * RD + (y + RB)) * RD + (x + RB); }
inline openvdb::CoordBBox rbox() { return openvdb::CoordBBox(openvdb::Coord(-RB), openvdb::Coord(RB)); }

/* The grids of main() :1157-1246, created the way main() creates them. */
struct World {
  openvdb::math::Transform::Ptr trans;
  openvdb::FloatGrid::Ptr container, output, solid, Adiag, Aplusj, Aplusi, Aplusk, diver, rhs;
  openvdb::Int32Grid::Ptr indices;
  openvdb::Vec3dGrid::Ptr vels, velBefore, normals;
  openvdb::BoolGrid::Ptr defined;
  PointList pos;
  Eigen::SparseMatrix<double> A;
  Eigen::VectorXd b, b2, p;
  Eigen::ConjugateGradient<Eigen::SparseMatrix<double>, Eigen::Lower | Eigen::Upper, Eigen::IncompleteCholesky<double>> cg;
  int numActive = 0;

  World() {
    trans = openvdb::math::Transform::createLinearTransform(1);
    auto mk = [&](float bg) { auto g = openvdb::FloatGrid::create(bg); g->setTransform(trans); g->fill(rbox(), 0, true); return g; };
    container = mk(0); output = mk(0); solid = mk(0);
    Adiag = mk(0); Aplusj = mk(0); Aplusi = mk(0); Aplusk = mk(0); diver = mk(0); rhs = mk(0);
    indices = openvdb::Int32Grid::create(0); indices->setTransform(trans);
    vels = openvdb::Vec3dGrid::create(); vels->setTransform(trans); vels->fill(rbox(), openvdb::Vec3d(0, 0, 0), true);
    velBefore = vels->deepCopy();
    normals = openvdb::Vec3dGrid::create(openvdb::Vec3d(0, 0, 0)); normals->fill(rbox(), openvdb::Vec3d(0, 0, 0), true);
    pos.initialize(60, trans);                                  /* :1347 */
    defined = pos.defined->deepCopy();                          /* :1362 */
  }
};

template <typename G, typename T>
void grid_to_array(G& g, T* out) {
  auto a = g->getAccessor();
  for (int z = -RB; z <= RB; ++z) for (int y = -RB; y <= RB; ++y) for (int x = -RB; x <= RB; ++x)
    out[ridx(x, y, z)] = (T)a.getValue(openvdb::Coord(x, y, z));
}
template <typename G, typename T>
void array_to_grid(G& g, const T* in) {
  auto a = g->getAccessor();
  typedef typename G::element_type::ValueType V;
  for (int z = -RB; z <= RB; ++z) for (int y = -RB; y <= RB; ++y) for (int x = -RB; x <= RB; ++x)
    a.setValue(openvdb::Coord(x, y, z), (V)in[ridx(x, y, z)]);
}
void vec_to_arrays(openvdb::Vec3dGrid::Ptr& g, double* u, double* v, double* w) {
  auto a = g->getAccessor();
  for (int z = -RB; z <= RB; ++z) for (int y = -RB; y <= RB; ++y) for (int x = -RB; x <= RB; ++x) {
    openvdb::Vec3d q = a.getValue(openvdb::Coord(x, y, z));
    size_t i = ridx(x, y, z); u[i] = q.x(); v[i] = q.y(); w[i] = q.z();
  }
}
void arrays_to_vec(openvdb::Vec3dGrid::Ptr& g, const double* u, const double* v, const double* w) {
  auto a = g->getAccessor();
  for (int z = -RB; z <= RB; ++z) for (int y = -RB; y <= RB; ++y) for (int x = -RB; x <= RB; ++x) {
    size_t i = ridx(x, y, z);
    a.setValue(openvdb::Coord(x, y, z), openvdb::Vec3d(u[i], v[i], w[i]));
  }
}

openvdb::FloatGrid::Ptr* fgrid(World* w, int which) {
  switch (which) {
    case 0: return &w->pos.weights; case 1: return &w->container; case 2: return &w->rhs; case 3: return &w->diver;
    case 4: return &w->Adiag; case 5: return &w->Aplusi; case 6: return &w->Aplusj; case 7: return &w->Aplusk;
    case 8: return &w->output; case 9: return &w->solid;
  }
  return nullptr;
}
openvdb::Vec3dGrid::Ptr* vgrid(World* w, int which) {
  switch (which) { case 0: return &w->vels; case 1: return &w->velBefore; case 2: return &w->normals; }
  return nullptr;
}

/* state of the last ref_main run */
struct MainRun {
  std::vector<double> pos0, pos, vel;            /* xyz interleaved */
  std::vector<double> u, v, w;
  std::vector<float> container, output, diver, rhs, adiag;
  std::vector<int32_t> indices;
  std::vector<long> passes;                      /* per step: Adiag->fill calls (one per do..while pass, :1461) */
  std::string log;
  int steps_done = 0;
} g_run;

}  // namespace

extern "C" {

int ref_grid_dim(void) { return RD; }
int ref_grid_origin(void) { return RB; }
double ref_spline(double x) { return spline(x); }
/* threads of the tbb::parallel_for stand-in (1 = serial, deterministic; > 1 for timing the reference arm) */
void ref_set_threads(int t) { openvdb::shim::state().threads = t < 1 ? 1 : t; }

void* ref_world_create(void) { return new World(); }
void ref_world_destroy(void* h) { delete static_cast<World*>(h); }

/* float grids: 0 weights (PointList member) 1 container 2 rhs 3 diver 4 Adiag 5 Aplusi 6 Aplusj 7 Aplusk 8 output 9 solid */
void ref_get_fgrid(void* h, int which, float* out) { grid_to_array(*fgrid(static_cast<World*>(h), which), out); }
void ref_set_fgrid(void* h, int which, const float* in) { array_to_grid(*fgrid(static_cast<World*>(h), which), in); }
/* Vec3d grids: 0 vels 1 velBeforeUpdate 2 normals */
void ref_get_vgrid(void* h, int which, double* u, double* v, double* w) { vec_to_arrays(*vgrid(static_cast<World*>(h), which), u, v, w); }
void ref_set_vgrid(void* h, int which, const double* u, const double* v, const double* w) { arrays_to_vec(*vgrid(static_cast<World*>(h), which), u, v, w); }
void ref_get_indices(void* h, int32_t* out) { grid_to_array(static_cast<World*>(h)->indices, out); }
void ref_set_indices(void* h, const int32_t* in, int numActive) {
  World* w = static_cast<World*>(h); array_to_grid(w->indices, in); w->numActive = numActive;
}
/* bool grids: 0 = main's `defined` (:1362), 1 = PointList::defined */
void ref_get_defined(void* h, int which, uint8_t* out) { World* w = static_cast<World*>(h); grid_to_array(which ? w->pos.defined : w->defined, out); }
void ref_set_defined(void* h, int which, const uint8_t* in) { World* w = static_cast<World*>(h); array_to_grid(which ? w->pos.defined : w->defined, in); }

void ref_set_particles(void* h, const double* px, const double* py, const double* pz, const double* vx, const double* vy,
                       const double* vz, int64_t n) {
  World* w = static_cast<World*>(h);
  w->pos.positions.clear(); w->pos.velocities.clear(); w->pos.initializedVelocities.clear();
  for (int64_t i = 0; i < n; ++i) {              /* direct stores: PointList::add (:841) would filter and zero the velocity */
    w->pos.positions.push_back(openvdb::Vec3d(px[i], py[i], pz[i]));
    w->pos.velocities.push_back(openvdb::Vec3d(vx[i], vy[i], vz[i]));
    w->pos.initializedVelocities.push_back(true);
  }
}
int64_t ref_num_particles(void* h) { return (int64_t) static_cast<World*>(h)->pos.size(); }
void ref_get_particles(void* h, double* px, double* py, double* pz, double* vx, double* vy, double* vz) {
  World* w = static_cast<World*>(h);
  for (size_t i = 0; i < w->pos.positions.size(); ++i) {
    px[i] = w->pos.positions[i].x(); py[i] = w->pos.positions[i].y(); pz[i] = w->pos.positions[i].z();
    vx[i] = w->pos.velocities[i].x(); vy[i] = w->pos.velocities[i].y(); vz[i] = w->pos.velocities[i].z();
  }
}
void ref_mark_uninitialized(void* h, int64_t first, int64_t count) {
  World* w = static_cast<World*>(h);
  for (int64_t i = first; i < first + count; ++i) w->pos.initializedVelocities[(size_t)i] = false;
}

/* ---- the reference's functions, called exactly as main() calls them ---- */
void ref_clear_vel(void* h) { World* w = static_cast<World*>(h); w->vels->fill(rbox(), openvdb::Vec3d(0, 0, 0), true); }         /* :1378 */
void ref_P2Gtransfer(void* h) { World* w = static_cast<World*>(h); w->pos.P2Gtransfer(w->vels, w->solid); }                       /* :1384 */
void ref_extrapolate(void* h) { World* w = static_cast<World*>(h); extrapolate(w->vels, w->pos.defined, w->pos.boundary, w->solid); } /* :1147 */
void ref_clear_container(void* h) { World* w = static_cast<World*>(h); w->container->fill(rbox(), 0, true); }                   /* :1396-1409 */
void ref_interpolate(void* h) { World* w = static_cast<World*>(h); w->pos.interpolate(w->container, w->solid); }                  /* :1413 */
void ref_save_vel_before(void* h) { World* w = static_cast<World*>(h); w->velBefore = w->vels->deepCopy(); }                      /* :1455 */
void ref_clear_pass_grids(void* h, int all) {                                                                                     /* :1461-1466 / :1477-1478 */
  World* w = static_cast<World*>(h);
  if (all) { w->Adiag->fill(rbox(), 0, true); w->Aplusj->fill(rbox(), 0, true); w->Aplusi->fill(rbox(), 0, true); w->Aplusk->fill(rbox(), 0, true); }
  w->diver->fill(rbox(), 0, true); w->rhs->fill(rbox(), 0, true);
}
void ref_setRHS(void* h, double dx, double gx, double gy, double gz, double dt) {
  World* w = static_cast<World*>(h); setRHS(dx, w->container, w->vels, w->rhs, w->solid, openvdb::Vec3d(gx, gy, gz), dt);         /* :1469 */
}
void ref_setDiver(void* h, double dx) { World* w = static_cast<World*>(h); setDiver(dx, w->vels, w->diver, w->rhs, w->container, w->solid); } /* :1470 */
void ref_setA(void* h, double dx, double dt, double rho) {
  World* w = static_cast<World*>(h); setA(dx, dt, rho, w->container, w->Adiag, w->Aplusi, w->Aplusj, w->Aplusk, w->solid);       /* :1471 */
}
/* setA2 (:1472): fills A and b; returns nnz.  A comes back in compressed column form. */
int64_t ref_setA2(void* h) {
  World* w = static_cast<World*>(h);
  w->A = Eigen::SparseMatrix<double>(w->numActive, w->numActive);                                                                /* :1460 */
  w->b = Eigen::VectorXd(w->numActive); w->b2 = Eigen::VectorXd(w->numActive);                                                   /* :1452-1453 */
  setA2(w->indices, w->b, w->A, w->diver, w->Adiag, w->Aplusi, w->Aplusj, w->Aplusk);
  return (int64_t)w->A.nonZeros();
}
void ref_get_A(void* h, int32_t* outer /* n+1 */, int32_t* inner /* nnz */, double* vals /* nnz */) {
  World* w = static_cast<World*>(h);
  w->A.makeCompressed();
  for (int i = 0; i <= w->A.outerSize(); ++i) outer[i] = w->A.outerIndexPtr()[i];
  for (int64_t k = 0; k < w->A.nonZeros(); ++k) { inner[k] = w->A.innerIndexPtr()[k]; vals[k] = w->A.valuePtr()[k]; }
}
void ref_get_b(void* h, int which, double* out) {
  World* w = static_cast<World*>(h); const Eigen::VectorXd& v = which == 0 ? w->b : which == 1 ? w->b2 : w->p;
  for (int i = 0; i < v.size(); ++i) out[i] = v[i];
}
int ref_solve(void* h, double* err) {                                                                                             /* :1473-1474 */
  World* w = static_cast<World*>(h);
  w->cg.compute(w->A);
  w->p = w->cg.solve(w->b);
  if (err) *err = w->cg.error();
  return (int)w->cg.iterations();
}
void ref_set_p(void* h, const double* p) { World* w = static_cast<World*>(h); w->p = Eigen::VectorXd(w->numActive); for (int i = 0; i < w->numActive; ++i) w->p[i] = p[i]; }
void ref_velUpdate(void* h, double dx, double dt_sub, double rho, double gx, double gy, double gz) {
  World* w = static_cast<World*>(h);
  velUpdate(w->indices, dx, dt_sub, rho, w->vels, w->container, w->p, w->solid, openvdb::Vec3d(gx, gy, gz), w->defined);         /* :1475 */
}
double ref_setOnlyB_error(void* h) {                                                                                              /* :1481-1483 */
  World* w = static_cast<World*>(h);
  setOnlyB(w->indices, w->b2, w->diver, w->Adiag);
  return ((w->b - w->b2).norm()) / (w->b.norm());
}
double ref_FLIPadvect(void* h, double maxDt, double dx, double dt) {
  World* w = static_cast<World*>(h);
  w->pos.FLIPadvect(maxDt, dx, dt, w->vels, 60, w->velBefore, w->container, w->solid, w->normals);                                /* :1490 */
  return dt;
}
double ref_advect(void* h, double maxDt, double dx, double dt) {
  World* w = static_cast<World*>(h);
  w->pos.advect(maxDt, dx, dt, w->vels, 60, w->container, w->solid);                                                              /* :1488 (commented out) */
  return dt;
}
void ref_interpFromGrid(void* h) { World* w = static_cast<World*>(h); w->pos.interpFromGrid(w->vels, 60, w->container); }        /* :1497 */

/* ---- the whole program: fluid.cc main() for `steps` steps of its own loop ---- */
int ref_main_run(int steps) {
  g_run = MainRun();
  auto& st = openvdb::shim::state();
  const size_t base = st.created.size();            /* grids that exist already (live Worlds) come first */
  st.point_lists.clear();
  long fills_before = 0;
  st.on_file_write = [&](const std::string& name) {
    if (name.compare(0, 11, "simulation/") != 0) return;       /* only the per-step file2.write, :1503 */
    /* creation order in main(): container, output, solid, fluid, Adiag, Aplusj, Aplusi, Aplusk, diver, rhs,
       indices, vels, normals (:1157-1246), then PointList::initialize -> weights, defined (:816-825) */
    auto G = [&](size_t k) { return st.created.at(base + k); };
    long fills = G(4)->fill_count;
    g_run.passes.push_back(fills - fills_before);
    fills_before = fills;
    PointList* pl = static_cast<PointList*>(st.point_lists.at(0));
    if (g_run.steps_done == 0) {
      /* nothing: pos0 is taken from the scatterer below */
    }
    ++g_run.steps_done;
    if (g_run.steps_done < steps) return;
    size_t n = pl->positions.size();
    g_run.pos.resize(3 * n); g_run.vel.resize(3 * n);
    for (size_t i = 0; i < n; ++i)
      for (int a = 0; a < 3; ++a) { g_run.pos[3 * i + a] = pl->positions[i][a]; g_run.vel[3 * i + a] = pl->velocities[i][a]; }
    auto F = [&](size_t k, std::vector<float>& out) {
      out.resize(RN); openvdb::FloatGrid* g = static_cast<openvdb::FloatGrid*>(G(k));
      for (int z = -RB; z <= RB; ++z) for (int y = -RB; y <= RB; ++y) for (int x = -RB; x <= RB; ++x)
        out[ridx(x, y, z)] = g->get(openvdb::Coord(x, y, z));
    };
    F(0, g_run.container); F(1, g_run.output); F(8, g_run.diver); F(9, g_run.rhs); F(4, g_run.adiag);
    g_run.indices.resize(RN);
    openvdb::Int32Grid* ig = static_cast<openvdb::Int32Grid*>(G(10));
    openvdb::Vec3dGrid* vg = static_cast<openvdb::Vec3dGrid*>(G(11));
    g_run.u.resize(RN); g_run.v.resize(RN); g_run.w.resize(RN);
    for (int z = -RB; z <= RB; ++z) for (int y = -RB; y <= RB; ++y) for (int x = -RB; x <= RB; ++x) {
      size_t i = ridx(x, y, z); openvdb::Coord c(x, y, z);
      g_run.indices[i] = ig->get(c);
      openvdb::Vec3d q = vg->get(c); g_run.u[i] = q.x(); g_run.v[i] = q.y(); g_run.w[i] = q.z();
    }
    throw openvdb::shim::StopRun();
  };
  std::ostringstream log;
  std::streambuf* old = std::cout.rdbuf(log.rdbuf());
  std::streamsize oldp = std::cout.precision(17);     /* so that "DT" / "Error:" (:1386,1491,1499) print every digit */
  int rc = 1;
  try { ref_main(0, nullptr); } catch (const openvdb::shim::StopRun&) { rc = 0; }
  std::cout.rdbuf(old); std::cout.precision(oldp);
  st.on_file_write = nullptr;
  g_run.log = log.str();
  /* initial particles = what the scatterer offered and PointList::add (:841) accepted */
  g_run.pos0.clear();
  for (const openvdb::Vec3d& p : st.last_scatter)
    if (abs(p.x()) < 60 - 2 && abs(p.y()) < 60 - 2 && abs(p.z()) < 60 - 2) { g_run.pos0.push_back(p.x()); g_run.pos0.push_back(p.y()); g_run.pos0.push_back(p.z()); }
  return rc;
}
int64_t ref_main_num_particles(void) { return (int64_t)(g_run.pos.size() / 3); }
int64_t ref_main_num_initial(void) { return (int64_t)(g_run.pos0.size() / 3); }
int ref_main_steps_done(void) { return g_run.steps_done; }
void ref_main_passes(long* out) { for (size_t i = 0; i < g_run.passes.size(); ++i) out[i] = g_run.passes[i]; }
void ref_main_particles(double* pos0, double* pos, double* vel) {
  if (pos0) std::memcpy(pos0, g_run.pos0.data(), g_run.pos0.size() * 8);
  if (pos) std::memcpy(pos, g_run.pos.data(), g_run.pos.size() * 8);
  if (vel) std::memcpy(vel, g_run.vel.data(), g_run.vel.size() * 8);
}
void ref_main_vels(double* u, double* v, double* w) {
  std::memcpy(u, g_run.u.data(), RN * 8); std::memcpy(v, g_run.v.data(), RN * 8); std::memcpy(w, g_run.w.data(), RN * 8);
}
/* 0 container 1 output 2 diver 3 rhs 4 Adiag */
void ref_main_fgrid(int which, float* out) {
  const std::vector<float>* s[] = {&g_run.container, &g_run.output, &g_run.diver, &g_run.rhs, &g_run.adiag};
  std::memcpy(out, s[which]->data(), RN * 4);
}
void ref_main_indices(int32_t* out) { std::memcpy(out, g_run.indices.data(), RN * 4); }
int64_t ref_main_log(char* out, int64_t cap) {
  int64_t n = (int64_t)g_run.log.size();
  if (out && cap > 0) { int64_t m = n < cap - 1 ? n : cap - 1; std::memcpy(out, g_run.log.data(), (size_t)m); out[m] = 0; }
  return n;
}

}  // extern "C"
