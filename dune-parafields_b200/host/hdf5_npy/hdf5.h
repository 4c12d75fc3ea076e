/* hdf5.h -- dependency-free stand-in for the handful of HDF5 calls behind
 * parafields' field files (io.hh:55-286: readParallelFromHDF5 /
 * writeParallelToHDF5, used by StochasticPart's file constructor and
 * writeToFile, stochastic.hh:80-107,273-294).
 *
 * HDF5 (and MPI-IO) are absent from this image.  Compiling the reference with
 * -DHAVE_HDF5=1 and this directory on the include path lets its OWN I/O layer
 * run unmodified -- writeToFile (".stoch.h5" + XDMF sidecar + ".trend" +
 * ".field" report), the file-reading constructors (randomfield.hh:190-225) --
 * on a container that needs no library:
 *
 *   <name>.stoch.h5  =  a NumPy .npy file (format 1.0): magic, a padded ASCII
 *   header {'descr': '<f8', 'fortran_order': False, 'shape': (n_{d-1},..,n_0)}
 *   and the values, last index fastest -- exactly the dimension order the
 *   reference gives HDF5 (io.hh:196-206: global_dim_HDF5[dim-i-1] =
 *   global_dim[i]).  `numpy.load(path)` reads it whatever the extension.
 *
 * It is NOT an HDF5 file: the XDMF sidecar the reference writes next to it
 * (Format="HDF") cannot be opened by HDF5 readers.  One dataset per file (the
 * reference writes exactly one, "/stochastic"); hyperslabs (count/offset per
 * dimension, no stride / block) as the reference selects them; several ranks
 * may write disjoint hyperslabs of one file (pwrite at computed offsets).
 * Only what io.hh calls is provided. */
#ifndef PF_HDF5_NPY_STANDIN_H
#define PF_HDF5_NPY_STANDIN_H

#include <cassert>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include <mpi.h>

typedef long long hid_t;
typedef int herr_t;
typedef unsigned long long hsize_t;

enum
{
  H5P_DEFAULT = 0,
  H5P_FILE_ACCESS = 1,
  H5P_DATASET_XFER = 2,
  H5F_ACC_RDONLY = 0,
  H5F_ACC_TRUNC = 2,
  H5S_SELECT_SET = 0,
  H5FD_MPIO_COLLECTIVE = 1,
  H5T_IEEE_F64LE = 64,
  H5T_IEEE_F32LE = 32,
  H5T_NATIVE_FLOAT = 132,
  H5T_NATIVE_DOUBLE = 164,
  H5T_NATIVE_LDOUBLE = 180
};

namespace pf_npy {

struct Object
{
  int kind = 0; // 1 property list, 2 file, 3 dataspace, 4 dataset
  // file
  std::string path;
  bool writable = false;
  // dataspace / dataset
  std::vector<hsize_t> dims;
  std::vector<hsize_t> offset, count; // selection (empty: everything)
  hid_t file = -1;
  long data_start = 0; // byte offset of the values in the file
};

inline std::map<hid_t, Object>&
table()
{
  static std::map<hid_t, Object> t;
  return t;
}
inline hid_t
add(const Object& o)
{
  static hid_t next = 1;
  table()[next] = o;
  return next++;
}
inline Object*
get(hid_t id, int kind)
{
  auto it = table().find(id);
  if (it == table().end() || it->second.kind != kind)
    return nullptr;
  return &it->second;
}

// .npy 1.0 header for a C-ordered array of little-endian doubles
inline std::string
header(const std::vector<hsize_t>& dims)
{
  std::string d = "{'descr': '<f8', 'fortran_order': False, 'shape': (";
  for (size_t i = 0; i < dims.size(); i++)
    d += std::to_string(dims[i]) + (dims.size() == 1 || i + 1 < dims.size() ? "," : "") +
         (i + 1 < dims.size() ? " " : "");
  d += "), }";
  const size_t unpadded = 10 + d.size() + 1;
  const size_t total = (unpadded + 63) / 64 * 64;
  d.append(total - unpadded, ' ');
  d += '\n';
  std::string h("\x93NUMPY\x01\x00", 8);
  const uint16_t len = (uint16_t)d.size();
  h += (char)(len & 0xff);
  h += (char)(len >> 8);
  return h + d;
}

// parse the header of an existing file: dims and the offset of the values
inline bool
parse(const std::string& path, std::vector<hsize_t>& dims, long& data_start)
{
  FILE* f = std::fopen(path.c_str(), "rb");
  if (!f)
    return false;
  unsigned char pre[10];
  bool ok = std::fread(pre, 1, 10, f) == 10 && std::memcmp(pre, "\x93NUMPY", 6) == 0 &&
            pre[6] == 1;
  std::string d;
  if (ok) {
    const size_t len = pre[8] + 256u * pre[9];
    d.resize(len);
    ok = std::fread(&d[0], 1, len, f) == len;
    data_start = 10 + (long)len;
  }
  std::fclose(f);
  if (!ok || d.find("'<f8'") == std::string::npos || d.find("False") == std::string::npos)
    return false;
  const size_t a = d.find("'shape': (");
  const size_t b = d.find(')', a);
  if (a == std::string::npos || b == std::string::npos)
    return false;
  dims.clear();
  const char* p = d.c_str() + a + 10;
  const char* end = d.c_str() + b;
  while (p < end) {
    char* q = nullptr;
    const unsigned long long v = std::strtoull(p, &q, 10);
    if (q == p)
      break;
    dims.push_back(v);
    p = q;
    while (p < end && (*p == ',' || *p == ' '))
      p++;
  }
  return !dims.empty();
}

// walk the selected hyperslab in file order: contiguous runs along the last
// dimension; `mem` is the dense local array in the same order
inline bool
transfer(Object& ds, const Object& space, void* mem, bool write)
{
  Object* file = get(ds.file, 2);
  if (!file)
    return false;
  const size_t rank = ds.dims.size();
  std::vector<hsize_t> off(rank, 0), cnt(ds.dims);
  if (!space.count.empty()) {
    off = space.offset;
    cnt = space.count;
  }
  hsize_t rows = 1;
  for (size_t i = 0; i + 1 < rank; i++)
    rows *= cnt[i];
  const hsize_t run = cnt[rank - 1];
  if (rows * run == 0)
    return true;
  FILE* f = std::fopen(file->path.c_str(), write ? "r+b" : "rb");
  if (!f)
    return false;
  std::vector<hsize_t> idx(rank, 0);
  char* m = static_cast<char*>(mem);
  bool ok = true;
  for (hsize_t r = 0; r < rows && ok; r++) {
    hsize_t lin = 0;
    for (size_t i = 0; i < rank; i++)
      lin = lin * ds.dims[i] + off[i] + (i + 1 < rank ? idx[i] : 0);
    ok = std::fseek(f, ds.data_start + (long)(lin * 8), SEEK_SET) == 0;
    if (ok)
      ok = (write ? std::fwrite(m, 8, run, f) : std::fread(m, 8, run, f)) == run;
    m += run * 8;
    for (size_t i = rank - 1; i-- > 0;) {
      if (++idx[i] < cnt[i])
        break;
      idx[i] = 0;
    }
  }
  std::fclose(f);
  return ok;
}

} // namespace pf_npy

inline hid_t
H5Pcreate(int)
{
  pf_npy::Object o;
  o.kind = 1;
  return pf_npy::add(o);
}
inline herr_t H5Pset_fapl_mpio(hid_t, MPI_Comm, MPI_Info) { return 0; }
inline herr_t H5Pset_dxpl_mpio(hid_t, int) { return 0; }
inline herr_t
H5Pclose(hid_t id)
{
  pf_npy::table().erase(id);
  return 0;
}

inline hid_t
H5Fcreate(const char* name, unsigned, hid_t, hid_t)
{
  pf_npy::Object o;
  o.kind = 2;
  o.path = name;
  o.writable = true;
  // rank 0 truncates; the header is written with the dataset
  int rank = 0;
  MPI_Comm_rank(MPI_COMM_WORLD, &rank);
  if (rank == 0) {
    FILE* f = std::fopen(name, "wb");
    if (!f)
      return -1;
    std::fclose(f);
  }
  MPI_Barrier(MPI_COMM_WORLD);
  return pf_npy::add(o);
}
inline hid_t
H5Fopen(const char* name, unsigned, hid_t)
{
  pf_npy::Object o;
  o.kind = 2;
  o.path = name;
  FILE* f = std::fopen(name, "rb");
  if (!f)
    return -1;
  std::fclose(f);
  return pf_npy::add(o);
}
inline herr_t
H5Fclose(hid_t id)
{
  pf_npy::table().erase(id);
  return 0;
}

inline hid_t
H5Screate_simple(int rank, const hsize_t* dims, const hsize_t*)
{
  pf_npy::Object o;
  o.kind = 3;
  o.dims.assign(dims, dims + rank);
  return pf_npy::add(o);
}
inline herr_t
H5Sclose(hid_t id)
{
  pf_npy::table().erase(id);
  return 0;
}
inline int
H5Sget_simple_extent_ndims(hid_t id)
{
  pf_npy::Object* s = pf_npy::get(id, 3);
  return s ? (int)s->dims.size() : -1;
}
inline int
H5Sget_simple_extent_dims(hid_t id, hsize_t* dims, hsize_t*)
{
  pf_npy::Object* s = pf_npy::get(id, 3);
  if (!s)
    return -1;
  for (size_t i = 0; i < s->dims.size(); i++)
    dims[i] = s->dims[i];
  return (int)s->dims.size();
}
inline herr_t
H5Sselect_hyperslab(hid_t id, int, const hsize_t* start, const hsize_t*,
                    const hsize_t* count, const hsize_t*)
{
  pf_npy::Object* s = pf_npy::get(id, 3);
  if (!s)
    return -1;
  s->offset.assign(start, start + s->dims.size());
  s->count.assign(count, count + s->dims.size());
  for (size_t i = 0; i < s->dims.size(); i++)
    if (s->offset[i] + s->count[i] > s->dims[i])
      return -1;
  return 0;
}

inline hid_t
H5Dcreate(hid_t file, const char*, int, hid_t space, hid_t, hid_t, hid_t)
{
  pf_npy::Object* f = pf_npy::get(file, 2);
  pf_npy::Object* s = pf_npy::get(space, 3);
  if (!f || !s)
    return -1;
  pf_npy::Object d;
  d.kind = 4;
  d.file = file;
  d.dims = s->dims;
  const std::string h = pf_npy::header(d.dims);
  d.data_start = (long)h.size();
  int rank = 0;
  MPI_Comm_rank(MPI_COMM_WORLD, &rank);
  if (rank == 0) {
    // header + the full extent, so that every rank can write its part in place
    FILE* fp = std::fopen(f->path.c_str(), "r+b");
    if (!fp)
      return -1;
    std::fwrite(h.data(), 1, h.size(), fp);
    hsize_t total = 1;
    for (hsize_t v : d.dims)
      total *= v;
    if (total > 0) {
      std::fseek(fp, d.data_start + (long)(total * 8) - 1, SEEK_SET);
      std::fputc(0, fp);
    }
    std::fclose(fp);
  }
  MPI_Barrier(MPI_COMM_WORLD);
  return pf_npy::add(d);
}
inline hid_t
H5Dopen(hid_t file, const char*, hid_t)
{
  pf_npy::Object* f = pf_npy::get(file, 2);
  if (!f)
    return -1;
  pf_npy::Object d;
  d.kind = 4;
  d.file = file;
  if (!pf_npy::parse(f->path, d.dims, d.data_start))
    return -1;
  return pf_npy::add(d);
}
inline hid_t
H5Dget_space(hid_t dset)
{
  pf_npy::Object* d = pf_npy::get(dset, 4);
  if (!d)
    return -1;
  pf_npy::Object s;
  s.kind = 3;
  s.dims = d->dims;
  return pf_npy::add(s);
}
inline herr_t
H5Dclose(hid_t id)
{
  pf_npy::table().erase(id);
  return 0;
}
// mem_type: only native doubles are stored (the CUDA path is fp64)
inline herr_t
H5Dwrite(hid_t dset, int mem_type, hid_t, hid_t filespace, hid_t, const void* buf)
{
  pf_npy::Object* d = pf_npy::get(dset, 4);
  pf_npy::Object* s = pf_npy::get(filespace, 3);
  if (!d || !s || mem_type != H5T_NATIVE_DOUBLE)
    return -1;
  return pf_npy::transfer(*d, *s, const_cast<void*>(buf), true) ? 0 : -1;
}
inline herr_t
H5Dread(hid_t dset, int mem_type, hid_t, hid_t filespace, hid_t, void* buf)
{
  pf_npy::Object* d = pf_npy::get(dset, 4);
  pf_npy::Object* s = pf_npy::get(filespace, 3);
  if (!d || !s || mem_type != H5T_NATIVE_DOUBLE)
    return -1;
  return pf_npy::transfer(*d, *s, buf, false) ? 0 : -1;
}

#endif /* PF_HDF5_NPY_STANDIN_H */
