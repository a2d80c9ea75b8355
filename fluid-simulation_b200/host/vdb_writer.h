// vdb_writer.h -- minimal, self-contained writer of OpenVDB files (file format 224, the format
// of the reference's vendored OpenVDB 4.0.2) for dense fp32 grids.  It stands in for
// openvdb::io::File::write(grids) at fluid.cc:1503,1508: no OpenVDB/TBB/Boost is available on the
// box, so `mygrids.vdb` is produced by this writer.  Layout follows, field by field:
//   header            openvdb/io/Archive.cc:939-971  (magic, 224, lib 4.0, hasGridOffsets, 36-char uuid)
//   file metadata     openvdb/MetaMap.cc:117         (empty map), then int32 grid count (Archive.cc:1166-1185)
//   per grid          openvdb/io/Archive.cc:1249-1327: descriptor (GridDescriptor.cc:81-100, unique
//                     name "name\x1eN" :178-182, type "Tree_float_5_4_3", instance parent, 3 x int64
//                     offsets), uint32 compression flags (Archive.cc:704-723), grid metadata,
//                     transform ("UniformScaleMap" + 5 Vec3d, math/Maps.h:843-850), topology
//                     (tree/Tree.h:1297,1439; tree/RootNode.h:2257-2290; tree/InternalNode.h:2175-2195;
//                     tree/LeafNode.h:1321), buffers (tree/LeafNode.h:1444-1452), values through
//                     io::writeCompressedValues (io/Compression.h:474-640) with COMPRESS_ACTIVE_MASK
//                     [| COMPRESS_ZIP], the reference build's default (Archive.cc:88-90, no Blosc).
#ifndef FLIPB200_VDB_WRITER_H
#define FLIPB200_VDB_WRITER_H
#include <stdint.h>
#include <stdio.h>
#include <set>
#include <string>

namespace flipvdb {

// Dense fp32 block of voxels [lo, lo+dim) in index space; data is x-fastest:
// idx = ((z-lo[2])*dim[1] + (y-lo[1]))*dim[0] + (x-lo[0]).  Every voxel of the block is written
// as ACTIVE with its value (the reference's outputGrid is fill()ed active over its whole box,
// fluid.cc:1161-1164); background is 0.
struct DenseGrid {
  int lo[3];
  int dim[3];
  const float* data;
  std::string name;      // fluid.cc never names its grids: ""
};

class Writer {
 public:
  // zip: COMPRESS_ZIP | COMPRESS_ACTIVE_MASK (the reference default) else COMPRESS_ACTIVE_MASK only
  Writer(const std::string& path, int grid_count, bool zip);
  ~Writer();
  bool ok() const { return f_ != nullptr && !failed_; }
  bool write_grid(const DenseGrid& g);   // call exactly grid_count times
  bool close();

 private:
  void put(const void* p, size_t n);
  void put_string(const std::string& s);
  void put_values(const float* v, size_t count);   // metadata byte + (zipped) active values
  FILE* f_;
  bool zip_;
  bool failed_;
  int written_, count_;
  std::set<std::string> names_;   // descriptor names written so far (unique-name rule of Archive.cc:1199-1210)
};

}  // namespace flipvdb
#endif
