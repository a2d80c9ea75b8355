/*
 * oracle/shim/openvdb/openvdb.h -- TEST INFRASTRUCTURE (see oracle/flip_oracle.h).
 *
 * A dense-array stand-in for the small OpenVDB / TBB surface that the reference's fluid.cc uses, so
 * that /root/reference/fluid.cc compiles UNMODIFIED with g++ (`-I oracle/shim -I /root/reference`,
 * see oracle/Makefile target `fluidref`) although the vendored OpenVDB cannot be built here
 * (libopenvdb.so.4.0.2 is a missing blob; TBB, Boost and Half are not installed -- SURVEY.md 0.2).
 * The compiled reference functions then pin the restatement in oracle/flip_oracle.cc bit for bit
 * (tests/test_ref_pin.py).  Nothing here is product code and nothing here is copied from OpenVDB:
 * every class below only reproduces the observable behaviour fluid.cc depends on, cited per item.
 *
 *  - Coord           openvdb/math/Coord.h:61  only (Int32,Int32,Int32) and explicit (Int32) exist, so
 *                    Coord(double,double,double) truncates toward zero (fluid.cc:1017-1025 relies on it)
 *  - Vec3d           openvdb/math/Vec3.h:224-230 (length), :265-340 (compound ops, component-wise,
 *                    division is a true division per component), :499-570 (binary ops via the compound ones)
 *  - Grid<T>         a VDB grid that fluid.cc only ever uses as a dense box: getValue outside every
 *                    written/filled voxel returns the background (openvdb/tree/ValueAccessor.h getValue);
 *                    FloatGrid stores float, so every setValue(double) rounds to fp32 (the implicit
 *                    double -> float conversion at the call, e.g. fluid.cc:292, :605)
 *  - fill(bbox,v,active), tree().voxelizeActiveTiles(), deepCopy(), setTransform: openvdb/Grid.h:1341,
 *                    :636; topology operations have no observable effect on a dense array
 *  - tbb::parallel_for(first,last,f): run serially in index order (one legal schedule; the reference's
 *                    own result depends on the schedule only through fp32/fp64 summation order)
 *  - tools::UniformPointScatter: openvdb/tools/PointScatter.h:143-185, :420-428 -- same count rule
 *                    (Index64(ppv*vol)*activeVoxels), same two engine copies, same libstdc++
 *                    distributions; voxels are visited in dense x,y,z order instead of VDB tree order,
 *                    so positions are distribution-equivalent, not identical, to a real OpenVDB build
 *  - io::File        write() calls a hook (the bridge snapshots state there and may stop the run)
 *
 * Box: indices [-SHIM_LIM, SHIM_LIM]^3 are stored densely; a write outside throws (the reference never
 * does that on valid input: every write is clamped or guarded, fluid.cc:271-276, :644-663).
 */
#ifndef FLIPB200_ORACLE_SHIM_OPENVDB_H
#define FLIPB200_ORACLE_SHIM_OPENVDB_H

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <functional>
#include <iostream>
#include <memory>
#include <random>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#ifndef SHIM_LIM
#define SHIM_LIM 63
#endif

namespace openvdb {

typedef int32_t Int32;
typedef uint64_t Index64;

inline void initialize() {}

class Coord {
 public:
  Coord() { v[0] = v[1] = v[2] = 0; }
  explicit Coord(Int32 xyz) { v[0] = v[1] = v[2] = xyz; }
  Coord(Int32 x, Int32 y, Int32 z) { v[0] = x; v[1] = y; v[2] = z; }
  Int32 x() const { return v[0]; }
  Int32 y() const { return v[1]; }
  Int32 z() const { return v[2]; }
  Int32& x() { return v[0]; }
  Int32& y() { return v[1]; }
  Int32& z() { return v[2]; }
  Int32 operator[](size_t i) const { return v[i]; }
 private:
  Int32 v[3];
};

class CoordBBox {
 public:
  CoordBBox() {}
  CoordBBox(const Coord& lo, const Coord& hi) : mMin(lo), mMax(hi) {}
  const Coord& min() const { return mMin; }
  const Coord& max() const { return mMax; }
 private:
  Coord mMin, mMax;
};

class Vec3d {
 public:
  Vec3d() { mm[0] = mm[1] = mm[2] = 0.0; }
  Vec3d(double x, double y, double z) { mm[0] = x; mm[1] = y; mm[2] = z; }
  double& x() { return mm[0]; }
  double& y() { return mm[1]; }
  double& z() { return mm[2]; }
  double x() const { return mm[0]; }
  double y() const { return mm[1]; }
  double z() const { return mm[2]; }
  double operator[](size_t i) const { return mm[i]; }
  double& operator[](size_t i) { return mm[i]; }
  double length() const { return std::sqrt(double(mm[0] * mm[0] + mm[1] * mm[1] + mm[2] * mm[2])); }
  const Vec3d& operator+=(const Vec3d& o) { mm[0] += o.mm[0]; mm[1] += o.mm[1]; mm[2] += o.mm[2]; return *this; }
  const Vec3d& operator-=(const Vec3d& o) { mm[0] -= o.mm[0]; mm[1] -= o.mm[1]; mm[2] -= o.mm[2]; return *this; }
  template <typename S> const Vec3d& operator*=(S s) { mm[0] = mm[0] * s; mm[1] = mm[1] * s; mm[2] = mm[2] * s; return *this; }
  template <typename S> const Vec3d& operator/=(S s) { mm[0] /= s; mm[1] /= s; mm[2] /= s; return *this; }
 private:
  double mm[3];
};
typedef Vec3d Vec3R;
inline Vec3d operator+(const Vec3d& a, const Vec3d& b) { Vec3d r(a); r += b; return r; }
inline Vec3d operator-(const Vec3d& a, const Vec3d& b) { Vec3d r(a); r -= b; return r; }
template <typename S> inline Vec3d operator*(const Vec3d& a, S s) { Vec3d r(a); r *= s; return r; }
template <typename S> inline Vec3d operator*(S s, const Vec3d& a) { return a * s; }
template <typename S> inline Vec3d operator/(const Vec3d& a, S s) { Vec3d r(a); r /= s; return r; }

namespace math {
class Transform {
 public:
  typedef std::shared_ptr<Transform> Ptr;
  static Ptr createLinearTransform(double voxelSize = 1.0) { Ptr t(new Transform); t->mSize = voxelSize; return t; }
  double voxelSize() const { return mSize; }
 private:
  double mSize = 1.0;
};
}  // namespace math

class GridBase;

/* instrumentation the bridge (oracle/ref_bridge.cc) reads: grids in creation order, PointLists seen by
 * the scatterer, hook fired by io::File::write */
namespace shim {
struct State {
  std::vector<GridBase*> created;                 /* Grid<T>::create() order; deepCopy()s are not listed */
  std::vector<void*> point_lists;                 /* the PointAccessor handed to UniformPointScatter */
  std::vector<Vec3d> last_scatter;                /* points offered to PointList::add by the last operator() */
  std::function<void(const std::string&)> on_file_write;
  int threads = 1;                                /* tbb::parallel_for stand-in: 1 = serial in index order (parity runs) */
};
inline State& state() { static State s; return s; }
struct StopRun {};                                /* thrown by the hook to leave ref_main after K steps */
}  // namespace shim

class GridBase {
 public:
  typedef std::shared_ptr<GridBase> Ptr;
  virtual ~GridBase() {
    auto& c = shim::state().created;
    c.erase(std::remove(c.begin(), c.end(), this), c.end());
  }
  void setTransform(math::Transform::Ptr t) { mTransform = t; }
  Vec3d voxelSize() const { double s = mTransform ? mTransform->voxelSize() : 1.0; return Vec3d(s, s, s); }
  Vec3d indexToWorld(const Vec3d& p) const { double s = mTransform ? mTransform->voxelSize() : 1.0; return p * s; }
  long fill_count = 0;                            /* number of fill() calls (Adiag: one per relaxation pass, fluid.cc:1461) */
 protected:
  math::Transform::Ptr mTransform;
};
typedef std::vector<GridBase::Ptr> GridPtrVec;

template <typename T> struct ShimStore { typedef T type; };
template <> struct ShimStore<bool> { typedef unsigned char type; };

template <typename T>
class Grid : public GridBase {
 public:
  typedef std::shared_ptr<Grid> Ptr;
  typedef T ValueType;
  typedef typename ShimStore<T>::type StoreT;
  enum { L = SHIM_LIM, D = 2 * SHIM_LIM + 1 };

  struct TreeStub { void voxelizeActiveTiles() {} };

  class Accessor {
   public:
    explicit Accessor(Grid* g) : g(g) {}
    T getValue(const Coord& c) const { return g->get(c); }
    void setValue(const Coord& c, const T& v) { g->set(c, v); }
   private:
    Grid* g;
  };

  static Ptr create() { return create(T()); }
  static Ptr create(const T& background) {
    Ptr p(new Grid(background));
    shim::state().created.push_back(p.get());
    return p;
  }
  Accessor getAccessor() { return Accessor(this); }
  TreeStub& tree() { return mTree; }
  void fill(const CoordBBox& b, const T& value, bool active = true) {
    ++fill_count;
    for (int x = b.min().x(); x <= b.max().x(); ++x)
      for (int y = b.min().y(); y <= b.max().y(); ++y)
        for (int z = b.min().z(); z <= b.max().z(); ++z) {
          size_t i = index(Coord(x, y, z), true);
          mData[i] = value;
          mActive[i] = active ? 1 : 0;
        }
  }
  Ptr deepCopy() const { return Ptr(new Grid(*this)); }
  Index64 activeVoxelCount() const { Index64 n = 0; for (unsigned char a : mActive) n += a; return n; }
  bool isActive(const Coord& c) const { long i = index(c, false); return i >= 0 && mActive[(size_t)i]; }

  T get(const Coord& c) const { long i = index(c, false); return i < 0 ? mBackground : T(mData[(size_t)i]); }
  void set(const Coord& c, const T& v) { size_t i = (size_t)index(c, true); mData[i] = v; mActive[i] = 1; }
  const T& background() const { return mBackground; }

 private:
  explicit Grid(const T& bg) : mBackground(bg), mData((size_t)D * D * D, StoreT(bg)), mActive((size_t)D * D * D, 0) {}
  long index(const Coord& c, bool must) const {
    if (c.x() < -L || c.x() > L || c.y() < -L || c.y() > L || c.z() < -L || c.z() > L) {
      if (must) throw std::out_of_range("oracle/shim: grid write outside the dense box");
      return -1;
    }
    return ((long)(c.x() + L) * D + (c.y() + L)) * D + (c.z() + L);
  }
  T mBackground;
  std::vector<StoreT> mData;
  std::vector<unsigned char> mActive;
  TreeStub mTree;
};

typedef Grid<float> FloatGrid;
typedef Grid<Int32> Int32Grid;
typedef Grid<bool> BoolGrid;
typedef Grid<Vec3d> Vec3dGrid;

namespace io {
class File {
 public:
  explicit File(const std::string& name) : mName(name) {}
  void write(const GridPtrVec&) { if (shim::state().on_file_write) shim::state().on_file_write(mName); }
  void close() {}
 private:
  std::string mName;
};
}  // namespace io

namespace util { class NullInterrupter {}; }

namespace tools {
template <typename PointAccessorType, typename RandomGenerator, typename InterruptType = util::NullInterrupter>
class UniformPointScatter {
 public:
  UniformPointScatter(PointAccessorType& points, float pointsPerVolume, RandomGenerator& randGen)
      : mPoints(points), mPointsPerVolume(pointsPerVolume), mEngine(randGen) {
    shim::state().point_lists.push_back(static_cast<void*>(&points));
  }
  template <typename GridT>
  bool operator()(const GridT& grid) {
    const Index64 voxels = grid.activeVoxelCount();
    if (voxels == 0) return false;
    const Vec3d dim = grid.voxelSize();
    const Index64 target = Index64(mPointsPerVolume * dim[0] * dim[1] * dim[2]) * voxels;   /* PointScatter.h:150 */
    std::vector<Index64> list((size_t)target);
    RandomGenerator idEngine(mEngine);                      /* RandInt copies Rand01's engine copy (:159) */
    std::uniform_int_distribution<Index64> pick(0, voxels - 1);
    for (Index64 i = 0; i < target; ++i) list[(size_t)i] = pick(idEngine);
    std::sort(list.begin(), list.end());
    std::uniform_real_distribution<double> rand01;
    shim::state().last_scatter.clear();
    Index64 id = 0, next = 0;
    const int L = GridT::L;
    for (int x = -L; x <= L && next < target; ++x)
      for (int y = -L; y <= L && next < target; ++y)
        for (int z = -L; z <= L && next < target; ++z) {
          if (!grid.isActive(Coord(x, y, z))) continue;
          while (next < target && list[(size_t)next] == id) {
            const double rx = 0.5 + 1.0 * (rand01(mEngine) - 0.5);            /* getRand(), spread 1 (:400) */
            const double ry = 0.5 + 1.0 * (rand01(mEngine) - 0.5);
            const double rz = 0.5 + 1.0 * (rand01(mEngine) - 0.5);
            const Vec3d pos((x - 0.5) + rx, (y - 0.5) + ry, (z - 0.5) + rz);   /* getCoord() - offset + rand (:420-428) */
            const Vec3d w = grid.indexToWorld(pos);
            shim::state().last_scatter.push_back(w);
            mPoints.add(w);
            ++next;
          }
          ++id;
        }
    return true;
  }
 private:
  PointAccessorType& mPoints;
  float mPointsPerVolume;
  RandomGenerator mEngine;                                    /* Rand01 holds a COPY of the caller's engine (Math.h:147) */
};
}  // namespace tools

}  // namespace openvdb

namespace tbb {
/* threads == 1: serial, index order (deterministic; what the pinning tests use).  threads > 1 (timing runs of
 * the reference arm only): contiguous chunks on std::threads -- the reference's own per-node mutexes
 * (fluid.cc:290-294, :872-874) and its maxSpeed mutex (:984-989) are what make its bodies safe under TBB,
 * and they are just as safe here; chunk boundaries are multiples of 64 so that the vector<bool> bodies
 * (:899-902) never share a word. */
template <typename I0, typename I1, typename F>
inline void parallel_for(I0 first, I1 last, const F& f) {
  const size_t a = (size_t)first, b = (size_t)last;
  const int T = openvdb::shim::state().threads;
  if (T <= 1 || b - a < 4096) { for (size_t i = a; i < b; ++i) f(i); return; }
  const size_t chunk = (((b - a) + T - 1) / T + 63) & ~(size_t)63;
  std::vector<std::thread> pool;
  for (int t = 0; t < T; ++t) {
    const size_t lo = a + (size_t)t * chunk, hi = std::min(b, lo + chunk);
    if (lo >= hi) break;
    pool.emplace_back([lo, hi, &f]() { for (size_t i = lo; i < hi; ++i) f(i); });
  }
  for (std::thread& th : pool) th.join();
}
}  // namespace tbb

#endif
