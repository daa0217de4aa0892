// icon_rt_b200 -- command-line host of the B200 path, a drop-in for the reference's non-interactive
// `icon_rt` (icon_rt/hostCode.cu:703-968 main + common/pipeline.cu:224-253 Pipeline flags):
//
//   icon_rt_b200 file.ic [--num-cells N] [--lat-range a:b] [--lon-range a:b]        (hostCode.cu:106-129)
//                [--bgcolor r g b] [--sample-limit n] [--xf file] [--size w h]       (pipeline.cu:224-253)
//                [-fovy deg] [--camera vp.xyz vi.xyz vu.xyz]
//   extras:      [--gen n k levels] synthetic RnBk grid instead of a file, [--dt meters], [--jitter],
//                [--sharp] the high-opacity test LUT, [-o out.png], [--integrator ...] [--sampler ...],
//                [--gpus N] sort-first over N GPUs (icb_group_*; marcher), [--save-xf f.xf] [--save-ic f.ic]
//
// Everything device-side goes through the C ABI (include/icon_b200.h); this file needs no CUDA headers.
// Output: <name>.png (rows flipped like stbi_flip_vertically_on_write, pipeline.cu:733-738) and the
// reference's "FPS:" line (pipeline.cu:739) computed from the device time of the render kernel.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "icon_b200.h"
#include "icon_host.hpp"

using namespace icon_b200;

namespace {

// ---- minimal PNG writer (stored deflate blocks; no compression library needed) -------------------
uint32_t crcTable[256];
void crcInit()
{
  for (uint32_t n = 0; n < 256; ++n) {
    uint32_t c = n;
    for (int k = 0; k < 8; ++k) c = (c & 1) ? 0xedb88320u ^ (c >> 1) : c >> 1;
    crcTable[n] = c;
  }
}
uint32_t crcUpdate(uint32_t c, const uint8_t *p, size_t n)
{
  for (size_t i = 0; i < n; ++i) c = crcTable[(c ^ p[i]) & 0xff] ^ (c >> 8);
  return c;
}
void be32(std::vector<uint8_t> &v, uint32_t x) { for (int s = 24; s >= 0; s -= 8) v.push_back((uint8_t)(x >> s)); }
void chunk(FILE *f, const char *type, const std::vector<uint8_t> &data)
{
  std::vector<uint8_t> head;
  be32(head, (uint32_t)data.size());
  fwrite(head.data(), 1, 4, f);
  uint32_t c = crcUpdate(0xffffffffu, (const uint8_t *)type, 4);
  c = crcUpdate(c, data.data(), data.size()) ^ 0xffffffffu;
  fwrite(type, 1, 4, f);
  if (!data.empty()) fwrite(data.data(), 1, data.size(), f);
  std::vector<uint8_t> tail;
  be32(tail, c);
  fwrite(tail.data(), 1, 4, f);
}
bool writePNG(const std::string &path, const uint32_t *rgba, int w, int h, bool flip)
{
  FILE *f = fopen(path.c_str(), "wb");
  if (!f) return false;
  crcInit();
  static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
  fwrite(sig, 1, 8, f);
  std::vector<uint8_t> ihdr;
  be32(ihdr, (uint32_t)w); be32(ihdr, (uint32_t)h);
  ihdr.insert(ihdr.end(), {8, 6, 0, 0, 0});
  chunk(f, "IHDR", ihdr);
  std::vector<uint8_t> raw;
  raw.reserve((size_t)h * (w * 4 + 1));
  for (int y = 0; y < h; ++y) {
    const uint32_t *row = rgba + (size_t)(flip ? h - 1 - y : y) * w;
    raw.push_back(0);
    const uint8_t *b = (const uint8_t *)row;
    raw.insert(raw.end(), b, b + (size_t)w * 4);
  }
  std::vector<uint8_t> z = {0x78, 0x01};
  uint32_t a = 1, b2 = 0;
  for (size_t i = 0; i < raw.size(); ++i) { a = (a + raw[i]) % 65521u; b2 = (b2 + a) % 65521u; }
  for (size_t off = 0; off < raw.size() || off == 0; off += 65535) {
    const size_t n = std::min<size_t>(65535, raw.size() - off);
    z.push_back(off + n >= raw.size() ? 1 : 0);
    z.push_back((uint8_t)(n & 0xff)); z.push_back((uint8_t)(n >> 8));
    z.push_back((uint8_t)(~n & 0xff)); z.push_back((uint8_t)((~n >> 8) & 0xff));
    z.insert(z.end(), raw.begin() + off, raw.begin() + off + n);
    if (raw.empty()) break;
  }
  be32(z, (b2 << 16) | a);
  chunk(f, "IDAT", z);
  chunk(f, "IEND", {});
  fclose(f);
  return true;
}

bool endsWith(const std::string &s, const std::string &suf)
{
  return s.size() >= suf.size() && s.compare(s.size() - suf.size(), suf.size(), suf) == 0;
}

void usage(FILE *f)
{
  fprintf(f,
          "Usage: icon_rt_b200 file.ic | --gen n k levels   [options]\n"
          "  reference flags (hostCode.cu:106-129, pipeline.cu:224-253):\n"
          "    --num-cells N  --lat-range a:b  --lon-range a:b  -mode m (accepted, ignored)\n"
          "    --bgcolor r g b (viewer-only in the reference, ignored)  --sample-limit n  --xf file.xf\n"
          "    --size w h | -win w h  -fovy deg  --camera vp.xyz vi.xyz vu.xyz\n"
          "  extras:\n"
          "    --integrator march|woodcock|woodcock_grid|woodcock_sphere   --sampler cell|prism\n"
          "    --gen n k levels (synthetic RnBk grid)  --dt meters  --jitter  --sharp  --gpus N (marcher: sort-first over N GPUs)\n"
          "    --save-xf file.xf (the transfer function as used)  --save-ic file.ic (the records as loaded / generated)  -o out.png\n");
}

#define CHECK(call)                                                                         \
  do {                                                                                      \
    int rc_ = (call);                                                                       \
    if (rc_ < 0) {                                                                          \
      fprintf(stderr, "%s failed (%d): %s\n", #call, rc_, icb_last_error(ctx));             \
      return 1;                                                                             \
    }                                                                                       \
  } while (0)

}  // namespace

int main(int argc, char **argv)
{
  std::string filepath, xfFile, saveXf, saveIc, outName = "icon_rt.png";
  int gpus = 1;
  long maxNumCells = -1;
  float latLo = -90.f, latHi = 90.f, lonLo = -180.f, lonHi = 180.f;
  float bg[3] = {0.f, 0.f, 0.f};
  int sampleLimit = 1, width = 512, height = 512;    // Frame fb(512,512), hostCode.cu:815
  float fovyDeg = 0.f, dt = 0.f;
  bool haveCamera = false, jitter = false, sharp = false, woodcock = false, prism = false;
  int integrator = ICB_INTEGRATOR_MARCH;
  vec3f vp{}, vi{}, vu{};
  int genN = 0, genK = 0, genL = 0;
  for (int i = 1; i < argc; ++i) {
    const std::string arg = argv[i];
    if (arg == "--help" || arg == "-h") { usage(stdout); return 0; }
    auto need = [&](int n) { if (i + n >= argc) { fprintf(stderr, "missing value after %s\n", arg.c_str()); exit(-1); } };
    if (arg[0] != '-' && endsWith(arg, ".ic")) filepath = arg;
    else if (arg == "--num-cells") { need(1); maxNumCells = std::atol(argv[++i]); }
    else if (arg == "--lat-range" || arg == "--lon-range") {
      need(1);
      const std::string s = argv[++i];
      const float lo = std::stof(s.substr(0, s.find(':'))), hi = std::stof(s.substr(s.find(':') + 1));
      if (arg == "--lat-range") { latLo = lo; latHi = hi; } else { lonLo = lo; lonHi = hi; }
    }
    else if (arg == "-mode") { need(1); ++i; }                   // sampler modes of the OWL build: accepted, ignored
    else if (arg == "--bgcolor") { need(3); for (int k = 0; k < 3; ++k) bg[k] = std::stof(argv[++i]); }
    else if (arg == "--sample-limit") { need(1); sampleLimit = std::atoi(argv[++i]); }
    else if (arg == "--xf") { need(1); xfFile = argv[++i]; }
    else if (arg == "-win" || arg == "--win" || arg == "--size") { need(2); width = std::atoi(argv[++i]); height = std::atoi(argv[++i]); }
    else if (arg == "-fovy") { need(1); fovyDeg = std::stof(argv[++i]); }
    else if (arg == "--camera") {
      need(9);
      vp = {std::stof(argv[i + 1]), std::stof(argv[i + 2]), std::stof(argv[i + 3])};
      vi = {std::stof(argv[i + 4]), std::stof(argv[i + 5]), std::stof(argv[i + 6])};
      vu = {std::stof(argv[i + 7]), std::stof(argv[i + 8]), std::stof(argv[i + 9])};
      i += 9;
      haveCamera = true;
    }
    else if (arg == "--gen") { need(3); genN = std::atoi(argv[++i]); genK = std::atoi(argv[++i]); genL = std::atoi(argv[++i]); }
    else if (arg == "--dt") { need(1); dt = std::stof(argv[++i]); }
    else if (arg == "--jitter") jitter = true;
    else if (arg == "--integrator") {     // march (default) | woodcock (woodcockTrackingAE) | woodcock_grid (woodcockTrackingWithAccel, the reference's default)
      need(1);
      const std::string v = argv[++i];
      if (v == "march") integrator = ICB_INTEGRATOR_MARCH;
      else if (v == "woodcock") integrator = ICB_INTEGRATOR_WOODCOCK;
      else if (v == "woodcock_grid") integrator = ICB_INTEGRATOR_WOODCOCK_GRID;
      else if (v == "woodcock_sphere") integrator = ICB_INTEGRATOR_WOODCOCK_SPHERE;
      else { fprintf(stderr, "unknown integrator %s\n", v.c_str()); return -1; }
      woodcock = integrator != ICB_INTEGRATOR_MARCH;
    }
    else if (arg == "--sampler") { need(1); prism = std::string(argv[++i]) == "prism"; }   // cell (default) | prism
    else if (arg == "--sharp") sharp = true;
    else if (arg == "-o") { need(1); outName = argv[++i]; }
    else if (arg == "--gpus") { need(1); gpus = std::atoi(argv[++i]); }
    else if (arg == "--save-xf") { need(1); saveXf = argv[++i]; }
    else if (arg == "--save-ic") { need(1); saveIc = argv[++i]; }
  }
  (void)bg;   // the marcher composites over transparent black like the reference's accumBuffer

  std::vector<ICONCell> cells;
  if (genN > 0) {
    const int64_t n = icb_gen_num_records(genN, genK, genL);
    if (n <= 0) { fprintf(stderr, "bad --gen arguments\n"); return -1; }
    cells.resize((size_t)n);
    if (icb_gen_grid(genN, genK, genL, 10.f, 0, cells.data(), n) != ICB_OK) { fprintf(stderr, "grid generation failed\n"); return -1; }
  } else {
    if (filepath.empty()) { usage(stderr); return -1; }
    if (!loadIC(filepath, cells, maxNumCells, latLo, latHi, lonLo, lonHi)) { fprintf(stderr, "cannot read %s\n", filepath.c_str()); return -1; }
  }
  printf("numCells: %zu\n", cells.size());
  if (!saveIc.empty()) {
    FILE *f = fopen(saveIc.c_str(), "wb");
    if (!f || fwrite(cells.data(), sizeof(ICONCell), cells.size(), f) != cells.size()) { fprintf(stderr, "cannot write %s\n", saveIc.c_str()); return 1; }
    fclose(f);
  }
  if (gpus < 1 || gpus > 64) { fprintf(stderr, "bad --gpus\n"); return -1; }
  if (gpus > 1 && integrator != ICB_INTEGRATOR_MARCH) { fprintf(stderr, "--gpus applies to the marcher\n"); return -1; }

  icb_handle ctx = nullptr;
  if (icb_create(0, &ctx) != ICB_OK) { fprintf(stderr, "%s\n", icb_last_error(nullptr)); return 1; }
  std::vector<icb_handle> handles{ctx};
  icb_group grp = nullptr;
  if (gpus > 1) {
    for (int d = 1; d < gpus; ++d) {
      icb_handle h = nullptr;
      if (icb_create(d, &h) != ICB_OK) { fprintf(stderr, "%s\n", icb_last_error(nullptr)); return 1; }
      handles.push_back(h);
    }
    if (icb_group_create_local(handles.data(), gpus, &grp) != ICB_OK) { fprintf(stderr, "%s\n", icb_group_last_error(nullptr)); return 1; }
    if (icb_group_set_grid(grp, cells.data(), cells.size()) != ICB_OK) { fprintf(stderr, "%s\n", icb_group_last_error(grp)); return 1; }
  } else
    CHECK(icb_set_grid(ctx, cells.data(), cells.size()));
  icb_bounds b;
  CHECK(icb_get_bounds(ctx, &b));

  Camera cam;
  cam.setAspect(width / (float)height);
  if (fovyDeg > 0.f) cam.fovy = deg2rad(fovyDeg);
  if (haveCamera) cam.setOrientation(vp, vi, vu, cam.fovy);
  else cam.viewAll(box3f{{b.volLo[0], b.volLo[1], b.volLo[2]}, {b.volHi[0], b.volHi[1], b.volHi[2]}});

  Transfunc tf;
  if (xfFile.empty() || !tf.loadXF(xfFile)) tf.setLUT(Transfunc::defaultLUT());
  tf.resampleIfSmall();
  if (sharp) for (int i = 0; i < tf.size(); ++i) tf.rgbaLUT[i].w = i < tf.size() / 2 ? 0.f : 1.f;
  if (!saveXf.empty() && !tf.saveXF(saveXf)) { fprintf(stderr, "cannot write %s\n", saveXf.c_str()); return 1; }
  for (icb_handle h : handles)
    if (icb_set_transfunc(h, &tf.rgbaLUT[0].x, tf.size(), tf.valueRange[0], tf.valueRange[1], tf.opacity) != ICB_OK) {
      fprintf(stderr, "icb_set_transfunc: %s\n", icb_last_error(h));
      return 1;
    }

  float org[3], d00[3], du[3], dv[3];
  cam.launchParams(width, height, org, d00, du, dv);
  for (icb_handle h : handles)
    if (icb_set_camera(h, org, d00, du, dv) != ICB_OK) { fprintf(stderr, "icb_set_camera: %s\n", icb_last_error(h)); return 1; }

  // the reference's own ray-gen (woodcockTrackingAE): one collision per pixel per frame, averaged over --sample-limit frames
  if (woodcock) CHECK(icb_set_integrator(ctx, integrator));
  if (prism)
    for (icb_handle h : handles)
      if (icb_set_sampler(h, ICB_SAMPLER_PRISM) != ICB_OK) { fprintf(stderr, "icb_set_sampler: %s\n", icb_last_error(h)); return 1; }

  const size_t npx = (size_t)width * height;
  void *dFb = nullptr, *dAccum = nullptr, *dDepth = nullptr;
  CHECK(icb_dev_alloc(ctx, npx * 4, &dFb));
  CHECK(icb_dev_alloc(ctx, npx * 16, &dAccum));
  CHECK(icb_dev_alloc(ctx, npx * 4, &dDepth));
  const float ambient[3] = {1.f, 1.f, 1.f};    // hostCode.cu:924-925
  float avg_t = 0.f;
  uint64_t steps = 0;
  std::vector<uint32_t> pixels(npx);
  if (grp) {
    // sort-first over the GPUs of this process: every rank renders its tiles, rank 0 assembles the frame (bit-identical
    // to the one-GPU frame) and it lands in page-locked host memory
    void *pinned = nullptr;
    CHECK(icb_host_alloc(ctx, npx * 4, &pinned));
    if (icb_group_set_params(grp, b.unitDistance, dt, ambient, 1.f, 0, jitter ? ICB_FLAG_JITTER : 0u, 0.f) != ICB_OK ||
        icb_group_begin(grp, width, height, 1) != ICB_OK) { fprintf(stderr, "%s\n", icb_group_last_error(grp)); return 1; }
    for (int frameID = 0; frameID < std::max(1, sampleLimit); ++frameID) {
      if (icb_group_render(grp, (uint32_t *)pinned) < 0 || icb_group_wait(grp, -1) != ICB_OK) { fprintf(stderr, "%s\n", icb_group_last_error(grp)); return 1; }
      float elapsed = 0.f;
      steps = 0;
      for (icb_handle h : handles) {
        icb_stats_t st;
        if (icb_stats(h, &st) != ICB_OK) return 1;
        elapsed = fmaxf(elapsed, st.lastKernelMs * 1e-3f);
        steps += st.steps;
      }
      avg_t = frameID == 0 ? elapsed : 0.8f * avg_t + 0.2f * elapsed;
    }
    std::memcpy(pixels.data(), pinned, npx * 4);
    icb_host_free(ctx, pinned);
  } else
  for (int frameID = 0; frameID < std::max(1, sampleLimit); ++frameID) {
    CHECK(icb_set_params(ctx, b.unitDistance, dt, ambient, 1.f, frameID, jitter ? ICB_FLAG_JITTER : 0u, 0.f));
    CHECK(icb_render(ctx, (uint32_t *)dFb, (float *)dAccum, (float *)dDepth, width, height, nullptr));
    icb_stats_t st;
    CHECK(icb_stats(ctx, &st));
    const float elapsed = st.lastKernelMs * 1e-3f;
    avg_t = frameID == 0 ? elapsed : 0.8f * avg_t + 0.2f * elapsed;   // pipeline.cu:590-606
    steps = st.steps;
  }
  if (!grp) CHECK(icb_dev_download(ctx, pixels.data(), dFb, npx * 4));
  if (!writePNG(outName, pixels.data(), width, height, true)) { fprintf(stderr, "cannot write %s\n", outName.c_str()); return 1; }
  printf("Output: %s\n", outName.c_str());
  printf("FPS: %.2f\n", 1.f / fmaxf(avg_t, 1e-8f));
  printf("samples/frame: %llu  samples/s: %.3e\n", (unsigned long long)steps, steps / fmaxf(avg_t, 1e-8f));
  icb_dev_free(ctx, dFb); icb_dev_free(ctx, dAccum); icb_dev_free(ctx, dDepth);
  if (grp) icb_group_destroy(grp);
  for (icb_handle h : handles) icb_destroy(h);
  return 0;
}
