// vdb_writer.cc -- see vdb_writer.h for the format map (file:line into the reference's OpenVDB).
#include "vdb_writer.h"
#include <string.h>
#include <stdlib.h>
#include <time.h>
#include <vector>
#include <random>
#include <zlib.h>

namespace flipvdb {

namespace {
const uint32_t COMPRESS_ZIP = 0x1, COMPRESS_ACTIVE_MASK = 0x2;     // io/Compression.h:78-83
const int8_t NO_MASK_OR_INACTIVE_VALS = 0;                         // io/Compression.h:93

inline int floor_div(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }

struct Buf {
  std::vector<char> d;
  void put(const void* p, size_t n) { const char* c = (const char*)p; d.insert(d.end(), c, c + n); }
  template <typename T> void pod(T v) { put(&v, sizeof(T)); }
};

inline void set_bit(std::vector<uint64_t>& w, int n) { w[n >> 6] |= (uint64_t)1 << (n & 63); }
}  // namespace

Writer::Writer(const std::string& path, int grid_count, bool zip)
    : f_(fopen(path.c_str(), "wb")), zip_(zip), failed_(false), written_(0), count_(grid_count) {
  if (!f_) return;
  // header, Archive.cc:939-971
  int64_t magic = 0x56444220;                    // OPENVDB_MAGIC, version.h:83
  put(&magic, 8);
  uint32_t v = 224; put(&v, 4);                  // OPENVDB_FILE_VERSION, version.h:96
  v = 4; put(&v, 4);                             // library major
  v = 0; put(&v, 4);                             // library minor
  char has_offsets = 1; put(&has_offsets, 1);    // seekable file
  std::mt19937 ran((uint32_t)(std::random_device()() + time(nullptr)));
  char uuid[37];
  const char* hex = "0123456789abcdef";
  for (int i = 0; i < 36; ++i) uuid[i] = (i == 8 || i == 13 || i == 18 || i == 23) ? '-' : hex[ran() & 15];
  uuid[14] = '4';                                // random (version 4) uuid, like boost's generator
  put(uuid, 36);
  uint32_t meta_count = 0; put(&meta_count, 4);  // file-level MetaMap is empty (File::write default)
  int32_t gc = grid_count; put(&gc, 4);
}

Writer::~Writer() { close(); }

void Writer::put(const void* p, size_t n) {
  if (!f_ || failed_) return;
  if (n && fwrite(p, 1, n, f_) != n) failed_ = true;
}

void Writer::put_string(const std::string& s) {   // util/Name.h:58-63
  uint32_t n = (uint32_t)s.size();
  put(&n, 4);
  put(s.data(), n);
}

bool Writer::close() {
  if (!f_) return false;
  bool ok = !failed_ && written_ == count_;
  if (fclose(f_) != 0) ok = false;
  f_ = nullptr;
  return ok;
}

// io::writeCompressedValues with COMPRESS_ACTIVE_MASK: all inactive values equal the background
// (0), so metadata = NO_MASK_OR_INACTIVE_VALS and only the `count` active values follow
// (io/Compression.h:515-612); with COMPRESS_ZIP they go through zipToStream (io/Compression.cc:69-98).
static void emit_values(Buf& out, const float* v, size_t count, bool zip) {
  out.pod<int8_t>(NO_MASK_OR_INACTIVE_VALS);
  const size_t bytes = count * sizeof(float);
  if (!zip) { out.put(v, bytes); return; }
  if (bytes == 0) { out.pod<int64_t>(0); return; }       // zipToStream of an empty buffer: size -0, no data
  uLongf zlen = compressBound((uLong)bytes);
  std::vector<Bytef> z(zlen);
  int st = compress2(z.data(), &zlen, (const Bytef*)v, (uLong)bytes, Z_DEFAULT_COMPRESSION);
  if (st == Z_OK && zlen < bytes) {
    out.pod<int64_t>((int64_t)zlen);
    out.put(z.data(), zlen);
  } else {
    out.pod<int64_t>(-(int64_t)bytes);
    out.put(v, bytes);
  }
}

bool Writer::write_grid(const DenseGrid& g) {
  if (!ok() || written_ >= count_) return false;
  const uint32_t flags = COMPRESS_ACTIVE_MASK | (zip_ ? COMPRESS_ZIP : 0);
  const int lo[3] = {g.lo[0], g.lo[1], g.lo[2]};
  const int hi[3] = {g.lo[0] + g.dim[0] - 1, g.lo[1] + g.dim[1] - 1, g.lo[2] + g.dim[2] - 1};
  const int64_t voxels = (int64_t)g.dim[0] * g.dim[1] * g.dim[2];

  // ---- tree: topology and buffers are generated in the same depth-first order ----
  Buf topo, bufs;
  int64_t leaf_count = 0;
  topo.pod<int32_t>(1);                                   // TreeBase::writeTopology: buffer count, Tree.h:1297-1301
  topo.pod<float>(0.0f);                                  // RootNode background, RootNode.h:2259-2265
  std::vector<int> ro;                                    // root children origins, x-major == Coord ordering
  for (int x = floor_div(lo[0], 4096) * 4096; x <= hi[0]; x += 4096)
    for (int y = floor_div(lo[1], 4096) * 4096; y <= hi[1]; y += 4096)
      for (int z = floor_div(lo[2], 4096) * 4096; z <= hi[2]; z += 4096) { ro.push_back(x); ro.push_back(y); ro.push_back(z); }
  topo.pod<uint32_t>(0);                                  // numTiles
  topo.pod<uint32_t>((uint32_t)(ro.size() / 3));          // numChildren
  auto overlaps = [&](const int o[3], int size) {
    return o[0] <= hi[0] && o[0] + size - 1 >= lo[0] && o[1] <= hi[1] && o[1] + size - 1 >= lo[1] &&
           o[2] <= hi[2] && o[2] + size - 1 >= lo[2];
  };
  std::vector<float> vals(512);
  for (size_t r = 0; r < ro.size(); r += 3) {
    const int O[3] = {ro[r], ro[r + 1], ro[r + 2]};
    topo.put(O, 12);                                      // child origin, RootNode.h:2284
    // InternalNode<.,5>: child mask, value mask, values (none active), then children (InternalNode.h:2175-2195)
    std::vector<uint64_t> cm2(512, 0), zero2(512, 0);
    std::vector<int> kids2;
    for (int n = 0; n < 32768; ++n) {
      const int b[3] = {O[0] + ((n >> 10) & 31) * 128, O[1] + ((n >> 5) & 31) * 128, O[2] + (n & 31) * 128};
      if (overlaps(b, 128)) { set_bit(cm2, n); kids2.push_back(n); }
    }
    topo.put(cm2.data(), 4096); topo.put(zero2.data(), 4096);
    emit_values(topo, nullptr, 0, zip_);
    for (int n : kids2) {
      const int B[3] = {O[0] + ((n >> 10) & 31) * 128, O[1] + ((n >> 5) & 31) * 128, O[2] + (n & 31) * 128};
      std::vector<uint64_t> cm1(64, 0), zero1(64, 0);
      std::vector<int> kids1;
      for (int m = 0; m < 4096; ++m) {
        const int l[3] = {B[0] + ((m >> 8) & 15) * 8, B[1] + ((m >> 4) & 15) * 8, B[2] + (m & 15) * 8};
        if (overlaps(l, 8)) { set_bit(cm1, m); kids1.push_back(m); }
      }
      topo.put(cm1.data(), 512); topo.put(zero1.data(), 512);
      emit_values(topo, nullptr, 0, zip_);
      for (int m : kids1) {
        const int L[3] = {B[0] + ((m >> 8) & 15) * 8, B[1] + ((m >> 4) & 15) * 8, B[2] + (m & 15) * 8};
        uint64_t mask[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        size_t cnt = 0;
        for (int bit = 0; bit < 512; ++bit) {             // leaf offset = x<<6 | y<<3 | z, LeafNode.h
          const int x = L[0] + (bit >> 6), y = L[1] + ((bit >> 3) & 7), z = L[2] + (bit & 7);
          if (x < lo[0] || x > hi[0] || y < lo[1] || y > hi[1] || z < lo[2] || z > hi[2]) continue;
          mask[bit >> 6] |= (uint64_t)1 << (bit & 63);
          vals[cnt++] = g.data[((size_t)(z - lo[2]) * g.dim[1] + (y - lo[1])) * g.dim[0] + (x - lo[0])];
        }
        topo.put(mask, 64);                               // LeafNode::writeTopology, LeafNode.h:1321
        bufs.put(mask, 64);                               // LeafNode::writeBuffers, LeafNode.h:1444-1452
        emit_values(bufs, vals.data(), cnt, zip_);
        ++leaf_count;
      }
    }
  }

  // ---- descriptor, Archive.cc:1249-1327 / GridDescriptor.cc:81-100 ----
  // unique descriptor name, Archive.cc:1199-1210: an EMPTY name always gets the suffix "\x1e0" (GridDescriptor::addSuffix,
  // GridDescriptor.cc:178-182), and any name already written gets "\x1e<n>" with the first free n >= 1.  (The reference
  // also suffixes the FIRST grid of a duplicated non-empty name, which a streaming writer cannot know; unnamed grids,
  // all fluid.cc writes, come out identically: "\x1e0", "\x1e1", ...)
  auto add_suffix = [](const std::string& nm, int n) { char b[32]; snprintf(b, sizeof(b), "\x1e%d", n); return nm + b; };
  std::string uname = g.name.empty() ? add_suffix(g.name, 0) : g.name;
  for (int n = 1; names_.count(uname); ++n) uname = add_suffix(g.name, n);
  names_.insert(uname);
  put_string(uname);
  put_string("Tree_float_5_4_3");                         // Tree::treeType(), tree/Tree.h:2129
  put_string("");                                         // instance parent
  const int64_t offset_pos = ftello(f_);
  int64_t pos[3] = {0, 0, 0};
  put(pos, 24);
  pos[0] = ftello(f_);                                    // gridPos
  put(&flags, 4);                                         // setGridCompression, Archive.cc:704-723
  // grid metadata (MetaMap.cc:117; the stats fields GridBase::addStatsMetadata adds, Grid.cc:446-457)
  uint32_t nmeta = 5; put(&nmeta, 4);
  auto meta_vec3i = [&](const char* name, const int v[3]) { put_string(name); put_string("vec3i"); uint32_t s = 12; put(&s, 4); put(v, 12); };
  auto meta_i64 = [&](const char* name, int64_t v) { put_string(name); put_string("int64"); uint32_t s = 8; put(&s, 4); put(&v, 8); };
  auto meta_str = [&](const char* name, const std::string& v) { put_string(name); put_string("string"); put_string(v); };
  meta_vec3i("file_bbox_max", hi);
  meta_vec3i("file_bbox_min", lo);
  meta_str("file_compression", zip_ ? "zip + active values" : "active values");
  meta_i64("file_mem_bytes", leaf_count * (int64_t)(512 * 4 + 96) + (int64_t)(ro.size() / 3) * (32768 * 8 + 4096 * 8));
  meta_i64("file_voxel_count", voxels);
  // transform: createLinearTransform(1) (fluid.cc:1154) is a UniformScaleMap; ScaleMap::write, Maps.h:843-850
  put_string("UniformScaleMap");
  const double one[3] = {1, 1, 1}, half[3] = {0.5, 0.5, 0.5};
  put(one, 24); put(one, 24); put(one, 24); put(one, 24); put(half, 24);
  put(topo.d.data(), topo.d.size());
  pos[1] = ftello(f_);                                    // blockPos
  put(bufs.d.data(), bufs.d.size());
  pos[2] = ftello(f_);                                    // endPos
  if (fseeko(f_, offset_pos, SEEK_SET) != 0) failed_ = true;
  put(pos, 24);
  if (fseeko(f_, pos[2], SEEK_SET) != 0) failed_ = true;
  ++written_;
  return !failed_;
}

}  // namespace flipvdb

// C entry points (tests / Python tooling): write `count` dense grids of identical box to one file; names[i] may be NULL / ""
extern "C" int flipvdb_write_file_named(const char* path, int count, const int lo[3], const int dim[3],
                                        const float* const* grids, int zip, const char* const* names) {
  flipvdb::Writer w(path, count, zip != 0);
  if (!w.ok()) return -1;
  for (int i = 0; i < count; ++i) {
    flipvdb::DenseGrid g;
    for (int a = 0; a < 3; ++a) { g.lo[a] = lo[a]; g.dim[a] = dim[a]; }
    if (names && names[i]) g.name = names[i];
    g.data = grids[i];
    if (!w.write_grid(g)) return -2;
  }
  return w.close() ? 0 : -3;
}
extern "C" int flipvdb_write_file(const char* path, int count, const int lo[3], const int dim[3],
                                  const float* const* grids, int zip) {
  return flipvdb_write_file_named(path, count, lo, dim, grids, zip, nullptr);
}
