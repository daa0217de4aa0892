// oracle/stub_nc/netcdf.h -- TEST INFRASTRUCTURE ONLY.
// A stand-in for the NetCDF-C calls tools/convert_icon/convert_icon.cpp makes (nc_open, nc_inq_dimid, nc_inq_dimlen,
// nc_inq_varid, nc_get_var_double, nc_get_var_int, nc_close, nc_strerror), so that the reference converter compiles and
// runs where it lies without the library: a "file" is a flat container of named dimensions and named double arrays
// (tools/gen_golden_convert.py writes them):   repeat { u8 kind (0 dim, 1 var) | u32 name length | name | u64 count |
// count doubles (var only) }.
// It also declares umesh::vec3f, which the converter's toCartesian uses outside its WITH_UMESH guard (convert_icon.cpp:48).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#define NC_NOWRITE 0
#define NC_NOERR 0

namespace umesh { struct vec3f { float x, y, z; }; }

namespace nc_stub {
struct File {
  std::vector<std::string> dimNames; std::vector<size_t> dimLen;
  std::vector<std::string> varNames; std::vector<std::vector<double>> vars;
};
inline std::map<int, File> &files() { static std::map<int, File> f; return f; }
inline int &nextId() { static int n = 1; return n; }
}

inline const char *nc_strerror(int e) { return e == NC_NOERR ? "no error" : "nc_stub: not found"; }

inline int nc_open(const char *path, int, int *ncid)
{
  FILE *f = std::fopen(path, "rb");
  if (!f) return -1;
  nc_stub::File F;
  for (;;) {
    unsigned char kind; uint32_t len; uint64_t count;
    if (std::fread(&kind, 1, 1, f) != 1) break;
    if (std::fread(&len, 4, 1, f) != 1) break;
    std::string name(len, '\0');
    if (len && std::fread(&name[0], 1, len, f) != len) break;
    if (std::fread(&count, 8, 1, f) != 1) break;
    if (kind == 0) { F.dimNames.push_back(name); F.dimLen.push_back((size_t)count); }
    else {
      std::vector<double> v((size_t)count);
      if (count && std::fread(v.data(), 8, (size_t)count, f) != (size_t)count) break;
      F.varNames.push_back(name); F.vars.push_back(std::move(v));
    }
  }
  std::fclose(f);
  *ncid = nc_stub::nextId()++;
  nc_stub::files()[*ncid] = std::move(F);
  return NC_NOERR;
}
inline int nc_close(int ncid) { nc_stub::files().erase(ncid); return NC_NOERR; }
inline int nc_inq_dimid(int ncid, const char *name, int *id)
{
  auto it = nc_stub::files().find(ncid);
  if (it == nc_stub::files().end()) return -1;
  for (size_t i = 0; i < it->second.dimNames.size(); ++i) if (it->second.dimNames[i] == name) { *id = (int)i; return NC_NOERR; }
  return -2;
}
inline int nc_inq_dimlen(int ncid, int id, size_t *len)
{
  auto it = nc_stub::files().find(ncid);
  if (it == nc_stub::files().end() || id < 0 || (size_t)id >= it->second.dimLen.size()) return -1;
  *len = it->second.dimLen[id];
  return NC_NOERR;
}
inline int nc_inq_varid(int ncid, const char *name, int *id)
{
  auto it = nc_stub::files().find(ncid);
  if (it == nc_stub::files().end()) return -1;
  for (size_t i = 0; i < it->second.varNames.size(); ++i) if (it->second.varNames[i] == name) { *id = (int)i; return NC_NOERR; }
  return -2;
}
inline int nc_get_var_double(int ncid, int id, double *out)
{
  auto it = nc_stub::files().find(ncid);
  if (it == nc_stub::files().end() || id < 0 || (size_t)id >= it->second.vars.size()) return -1;
  const std::vector<double> &v = it->second.vars[id];
  if (!v.empty()) std::memcpy(out, v.data(), v.size() * sizeof(double));
  return NC_NOERR;
}
inline int nc_get_var_int(int ncid, int id, int *out)
{
  auto it = nc_stub::files().find(ncid);
  if (it == nc_stub::files().end() || id < 0 || (size_t)id >= it->second.vars.size()) return -1;
  const std::vector<double> &v = it->second.vars[id];
  for (size_t i = 0; i < v.size(); ++i) out[i] = (int)v[i];
  return NC_NOERR;
}
