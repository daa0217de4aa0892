// C++ host mirror of the reference's host-side objects for the icon_rt path, sitting above the
// C ABI (include/icon_b200.h).  Same names, argument meaning and arithmetic as the reference so that
// a caller of icon_rt's host API finds what it expects:
//
//   Camera     common/camera.h:28-114   (setAspect, setOrientation, forceUpFrame, getScreen, viewAll)
//   Transfunc  common/transfunc.h:29-49 (opacity, valueRange, relRange, LUT) + resampleLUT
//              (common/dvr_course-common.h:44-70) + .xf files (common/pipeline.cu:127-169)
//   Frame      common/fb.h:27-48        (fbPointer, fbDepth, accumBuffer -- device memory, as under RTCORE)
//   ICONCell   icon_rt/ICONGrid.h:59-76 and the loader of icon_rt/hostCode.cu:717-758
//
// Compile with -ffp-contract=off (camera vectors feed the bit-exact ray generation).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

namespace icon_b200 {

struct vec3f {
  float x, y, z;
};
inline vec3f operator+(vec3f a, vec3f b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline vec3f operator-(vec3f a, vec3f b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline vec3f operator-(vec3f a) { return {-a.x, -a.y, -a.z}; }
inline vec3f operator*(vec3f a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline vec3f operator*(float s, vec3f a) { return {s * a.x, s * a.y, s * a.z}; }
inline vec3f operator/(vec3f a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline bool operator==(vec3f a, vec3f b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
inline float dot(vec3f a, vec3f b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline vec3f cross(vec3f u, vec3f v) { return {u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x}; }
inline float length(vec3f a) { return sqrtf(dot(a, a)); }
inline vec3f normalize(vec3f a) { return a / sqrtf(dot(a, a)); }

struct vec4f {
  float x, y, z, w;
};

struct box3f {
  vec3f lower, upper;
  vec3f center() const { return (lower + upper) / 2.f; }
  vec3f size() const { return upper - lower; }
};

#pragma pack(push, 1)
struct ICONCell {
  float   lat[3];
  float   lon[3];
  int32_t numLayers;
  float   height[32];
  float   value[32];
};
#pragma pack(pop)
static_assert(sizeof(ICONCell) == 284, "ICONCell must stay 284 bytes (.ic record)");

// ---- camera ----------------------------------------------------------------------------------
struct Camera {
  void setAspect(float a) { aspect = a; }

  void setOrientation(const vec3f &origin, const vec3f &poi, const vec3f &up, float fovyRadians)
  {
    position = origin;
    upVector = up;
    fovy = fovyRadians;
    frame.vz = (poi == origin) ? vec3f{0, 0, 1} : -normalize(poi - origin);   // looks down -z
    frame.vx = cross(up, frame.vz);
    if (dot(frame.vx, frame.vx) < 1e-8f) frame.vx = vec3f{0, 1, 0};
    else frame.vx = normalize(frame.vx);
    frame.vy = normalize(cross(frame.vz, frame.vx));
    distance = length(poi - origin);
    forceUpFrame();
  }

  void forceUpFrame()
  {
    if (fabsf(dot(frame.vz, upVector)) < 1e-6f) return;
    frame.vx = normalize(cross(upVector, frame.vz));
    frame.vy = normalize(cross(frame.vz, frame.vx));
  }

  vec3f getPosition() const { return position; }
  vec3f getPOI() const { return position - frame.vz * distance; }
  vec3f getUp() const { return upVector; }
  float getFovyInRadians() const { return fovy; }

  void getScreen(vec3f &lower_left, vec3f &horizontal, vec3f &vertical) const
  {
    const float screen_height = 2.f * tanf(0.5f * fovy);
    vertical = screen_height * frame.vy;
    horizontal = (screen_height * aspect) * frame.vx;
    lower_left = -frame.vz - 0.5f * vertical - 0.5f * horizontal;
  }

  void viewAll(const box3f &box)
  {
    const vec3f up{0, 1, 0};
    const float diagonal = length(box.size());
    const float r = diagonal * 0.5f;
    const vec3f eye = box.center() + vec3f{0, 0, r + r / std::atan(fovy)};
    setOrientation(eye, box.center(), up, fovy);
  }

  // camera.{org,dir_00,dir_du,dir_dv} as icon_rt/hostCode.cu:936-945 derives them
  void launchParams(int width, int height, float org[3], float d00[3], float du[3], float dv[3]) const
  {
    vec3f ll, hz, vt;
    getScreen(ll, hz, vt);
    const vec3f u = hz / (float)width, v = vt / (float)height;
    org[0] = position.x; org[1] = position.y; org[2] = position.z;
    d00[0] = ll.x; d00[1] = ll.y; d00[2] = ll.z;
    du[0] = u.x; du[1] = u.y; du[2] = u.z;
    dv[0] = v.x; dv[1] = v.y; dv[2] = v.z;
  }

  vec3f position{0, 0, 0}, upVector{0, 1, 0};
  float distance{0.f};
  float fovy{float(90.f * M_PI / 180.f)};
  float aspect{1.f};
  struct {
    vec3f vx{1, 0, 0}, vy{0, 1, 0}, vz{0, 0, 1};
  } frame;
};

// ---- transfer function ---------------------------------------------------------------------------
inline float lerpRef(float a, float b, float x) { return x * a + (1.f - x) * b; }   // vecmath.h:56-59: x weights a

inline void resampleLUT(std::vector<vec4f> &dst, const std::vector<vec4f> &src)
{
  const int srcDims = (int)src.size(), dstDims = (int)dst.size();
  for (int i = 0; i < dstDims; ++i) {
    const float indexf = i / (float)(dstDims) * (srcDims - 1);
    const int indexa = (int)indexf;
    const int indexb = indexa + 1 < srcDims - 1 ? indexa + 1 : srcDims - 1;
    const float frac = indexf - indexa;
    const vec4f &a = src[indexa], &b = src[indexb];
    dst[i] = {lerpRef(a.x, b.x, 1.f - frac), lerpRef(a.y, b.y, 1.f - frac), lerpRef(a.z, b.z, 1.f - frac),
              lerpRef(a.w, b.w, 1.f - frac)};
  }
}

struct Transfunc {
  float opacity{1.f};
  float valueRange[2]{0.f, 1.f};
  float relRange[2]{0.f, 1.f};
  std::vector<vec4f> rgbaLUT;

  void setLUT(const std::vector<vec4f> &lut) { rgbaLUT = lut; }
  const std::vector<vec4f> &getLUT() const { return rgbaLUT; }
  int size() const { return (int)rgbaLUT.size(); }

  // the 5-point default of icon_rt/hostCode.cu:828-834
  static std::vector<vec4f> defaultLUT()
  {
    return {{0.149f, 0.015f, 0.705f, 1.0f}, {0.486f, 0.603f, 0.956f, 0.75f}, {0.866f, 0.866f, 0.866f, 0.5f},
            {0.996f, 0.690f, 0.552f, 0.25f}, {0.752f, 0.298f, 0.231f, 0.0f}};
  }
  // non-interactive pipelines resample small LUTs to 300 entries (common/pipeline.cu:469-473)
  void resampleIfSmall()
  {
    if (size() < 300) {
      std::vector<vec4f> n(300);
      resampleLUT(n, rgbaLUT);
      rgbaLUT.swap(n);
    }
  }
  // `.xf` layout (common/pipeline.cu:127-169): opacity f32, valueRange 2xf32, relRange 2xf32, count i32, count x vec4f
  bool loadXF(const std::string &path)
  {
    std::ifstream in(path, std::ios::binary);
    if (!in.good()) return false;
    int32_t n = 0;
    in.read((char *)&opacity, 4);
    in.read((char *)valueRange, 8);
    in.read((char *)relRange, 8);
    in.read((char *)&n, 4);
    if (!in.good() || n < 0 || n > (1 << 24)) return false;
    rgbaLUT.resize((size_t)n);
    in.read((char *)rgbaLUT.data(), sizeof(vec4f) * (size_t)n);
    return in.good() || in.eof();
  }
  bool saveXF(const std::string &path) const
  {
    std::ofstream out(path, std::ios::binary);
    if (!out.good()) return false;
    const int32_t n = size();
    out.write((const char *)&opacity, 4);
    out.write((const char *)valueRange, 8);
    out.write((const char *)relRange, 8);
    out.write((const char *)&n, 4);
    out.write((const char *)rgbaLUT.data(), sizeof(vec4f) * (size_t)n);
    return out.good();
  }
};

// ---- grid loading -----------------------------------------------------------------------------------
inline float deg2rad(float d) { return d * float(M_PI) / 180.f; }   // ICONGrid.h:26-29

// icon_rt/hostCode.cu:717-758: numCells = filesize/284, optional --num-cells truncation, then every record
// with any corner outside the lat/lon window (degrees) is dropped; survivors keep their relative order.
inline bool loadIC(const std::string &path, std::vector<ICONCell> &cells, long maxNumCells = -1,
                   float latLo = -90.f, float latHi = 90.f, float lonLo = -180.f, float lonHi = 180.f)
{
  std::ifstream in(path, std::ios::binary);
  if (!in.good()) return false;
  in.seekg(0, in.end);
  size_t numCells = (size_t)in.tellg() / sizeof(ICONCell);
  in.seekg(0, in.beg);
  if (maxNumCells >= 0 && (size_t)maxNumCells < numCells) numCells = (size_t)maxNumCells;
  cells.resize(numCells);
  in.read((char *)cells.data(), sizeof(ICONCell) * numCells);
  const float a0 = deg2rad(latLo), a1 = deg2rad(latHi), o0 = deg2rad(lonLo), o1 = deg2rad(lonHi);
  size_t w = 0;
  for (size_t i = 0; i < cells.size(); ++i) {
    const ICONCell &c = cells[i];
    bool out = false;
    for (int k = 0; k < 3; ++k)
      out = out || c.lat[k] < a0 || c.lat[k] > a1 || c.lon[k] < o0 || c.lon[k] > o1;
    if (!out) {
      if (w != i) cells[w] = c;
      ++w;
    }
  }
  cells.resize(w);
  return true;
}

}  // namespace icon_b200
