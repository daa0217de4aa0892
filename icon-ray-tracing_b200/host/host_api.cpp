// C wrappers around the C++ host mirror (icon_host.hpp) so that Python/ctypes callers and tests get
// the same bits as a C++ caller (same libm, same operand order as the reference's camera.h etc.).
#include "icon_b200.h"
#include "icon_host.hpp"

using namespace icon_b200;

extern "C" {

void icb_camera_view_all(const float volLo[3], const float volHi[3], float fovy, float aspect, int width, int height,
                         float org[3], float dir_00[3], float dir_du[3], float dir_dv[3])
{
  Camera cam;
  cam.fovy = fovy;
  cam.setAspect(aspect);
  cam.viewAll(box3f{{volLo[0], volLo[1], volLo[2]}, {volHi[0], volHi[1], volHi[2]}});
  cam.launchParams(width, height, org, dir_00, dir_du, dir_dv);
}

void icb_camera_orient(const float eye[3], const float poi[3], const float up[3], float fovy, float aspect,
                       int width, int height, float org[3], float dir_00[3], float dir_du[3], float dir_dv[3])
{
  Camera cam;
  cam.setAspect(aspect);
  cam.setOrientation({eye[0], eye[1], eye[2]}, {poi[0], poi[1], poi[2]}, {up[0], up[1], up[2]}, fovy);
  cam.launchParams(width, height, org, dir_00, dir_du, dir_dv);
}

void icb_resample_lut(const float *src, int nsrc, float *dst, int ndst)
{
  std::vector<vec4f> s(nsrc), d(ndst);
  std::memcpy(s.data(), src, sizeof(vec4f) * (size_t)nsrc);
  resampleLUT(d, s);
  std::memcpy(dst, d.data(), sizeof(vec4f) * (size_t)ndst);
}

int64_t icb_load_ic(const char *path, int64_t maxNumCells, float latLo, float latHi, float lonLo, float lonHi,
                    void *records, int64_t capacity)
{
  if (!path) return ICB_ERR_INVALID;
  std::vector<ICONCell> cells;
  if (!loadIC(path, cells, (long)maxNumCells, latLo, latHi, lonLo, lonHi)) return ICB_ERR_INVALID;
  if (records) {
    if ((int64_t)cells.size() > capacity) return ICB_ERR_INVALID;
    std::memcpy(records, cells.data(), cells.size() * sizeof(ICONCell));
  }
  return (int64_t)cells.size();
}

}  // extern "C"
