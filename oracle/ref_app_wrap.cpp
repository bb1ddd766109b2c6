// ref_app_wrap.cpp — TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// Compiles the UNMODIFIED reference example application (examples/ultrahdr_app.cpp, included from
// where it lies under $(REF)) with its main() renamed, and exposes the one routine row N3 needs:
// UltraHdrAppInput::convertYuv420ToRGBImage (ultrahdr_app.cpp:682-734), the reference's own
// BT.601 YUV420 -> RGBA8888 arithmetic. The uhdr_* API functions the rest of that file calls are
// never reached; oracle/Makefile generates trap stubs for them so the library loads.
#define main uhdr_reference_app_main
#include "examples/ultrahdr_app.cpp"
#undef main

#include <fcntl.h>
#include <unistd.h>

extern "C" int ref_app_yuv420_to_rgba8888(const uint8_t* y, const uint8_t* u, const uint8_t* v, int w, int h,
                                          int y_stride, int u_stride, int v_stride, uint32_t* out) {
  UltraHdrAppInput app(nullptr, nullptr, nullptr, nullptr, nullptr, (size_t)w, (size_t)h);
  memset(&app.mRawYuv420Image, 0, sizeof app.mRawYuv420Image);
  memset(&app.mRawRgba8888Image, 0, sizeof app.mRawRgba8888Image);
  app.mRawYuv420Image.fmt = UHDR_IMG_FMT_12bppYCbCr420;
  app.mRawYuv420Image.w = (unsigned)w;
  app.mRawYuv420Image.h = (unsigned)h;
  app.mRawYuv420Image.planes[UHDR_PLANE_Y] = const_cast<uint8_t*>(y);
  app.mRawYuv420Image.planes[UHDR_PLANE_U] = const_cast<uint8_t*>(u);
  app.mRawYuv420Image.planes[UHDR_PLANE_V] = const_cast<uint8_t*>(v);
  app.mRawYuv420Image.stride[UHDR_PLANE_Y] = (unsigned)y_stride;
  app.mRawYuv420Image.stride[UHDR_PLANE_U] = (unsigned)u_stride;
  app.mRawYuv420Image.stride[UHDR_PLANE_V] = (unsigned)v_stride;
  // the routine also dumps its result to ./inRgba8888.raw: run it inside a scratch directory
  char tmpl[] = "/tmp/uhdr_ref_app_XXXXXX";
  const int back = open(".", O_RDONLY);
  const char* dir = mkdtemp(tmpl);
  if (back < 0 || !dir || chdir(dir) != 0) return -1;
  const bool ok = app.convertYuv420ToRGBImage();
  unlink("inRgba8888.raw");
  if (fchdir(back) != 0) return -1;
  close(back);
  rmdir(dir);
  if (ok) memcpy(out, app.mRawRgba8888Image.planes[UHDR_PLANE_PACKED], (size_t)w * h * 4);
  for (int i = 0; i < 3; i++) app.mRawYuv420Image.planes[i] = nullptr;  // caller's memory: keep the destructor off it
  return ok ? 0 : -1;
}
