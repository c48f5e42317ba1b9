// stitch_b200_napi.cc -- N-API addon over include/stitch_b200.h: what the TypeScript host of image-stitch loads
// (src/gpu/stitch-b200.ts in this repository shows the glue).  Thin on purpose: no pixel logic, only marshalling;
// every export maps to one C-ABI call.  Build (where Node headers exist; not in this repository's containers):
//   g++ -O2 -fPIC -shared -std=c++17 -I$NODE_INCLUDE -Iinclude native/stitch_b200_napi.cc (continued)
//       -Limage_stitch_b200 -lstitch_b200 -Wl,-rpath,'$ORIGIN' -o build/Release/stitch_b200_napi.node
// NOT compiled or run here (no node, no node_api.h: probed); kept in-tree so that a maintainer can.
#include <node_api.h>

#include <cstring>
#include <string>

#include "stitch_b200.h"

namespace {

#define NAPI_OK(call)                                        \
  do {                                                       \
    if ((call) != napi_ok) {                                 \
      napi_throw_error(env, nullptr, "N-API call failed: " #call); \
      return nullptr;                                        \
    }                                                        \
  } while (0)

struct Ctx { stitch_ctx *ctx = nullptr; };
struct Job { stitch_job *job = nullptr; Ctx *owner = nullptr; };

void FinalizeCtx(napi_env, void *data, void *) {
  auto *c = static_cast<Ctx *>(data);
  if (c->ctx) stitch_ctx_destroy(c->ctx);
  delete c;
}
void FinalizeJob(napi_env, void *data, void *) {
  auto *j = static_cast<Job *>(data);
  if (j->job) stitch_job_destroy(j->job);
  delete j;
}

// The message the reference's Error would carry (stitch_last_error of the job's OWN context).
napi_value ThrowStitch(napi_env env, Ctx *c, int rc) {
  const char *msg = c && c->ctx ? stitch_last_error(c->ctx) : nullptr;
  std::string text = msg && *msg ? msg : ("libstitch_b200 error " + std::to_string(rc));
  napi_throw_error(env, nullptr, text.c_str());
  return nullptr;
}

bool GetU32(napi_env env, napi_value obj, const char *key, uint32_t *out, uint32_t dflt = 0) {
  napi_value v;
  bool has = false;
  *out = dflt;
  if (napi_has_named_property(env, obj, key, &has) != napi_ok || !has) return true;
  if (napi_get_named_property(env, obj, key, &v) != napi_ok) return false;
  napi_valuetype t;
  if (napi_typeof(env, v, &t) != napi_ok) return false;
  if (t == napi_undefined || t == napi_null) return true;
  if (t == napi_boolean) { bool b; napi_get_value_bool(env, v, &b); *out = b ? 1u : 0u; return true; }
  return napi_get_value_uint32(env, v, out) == napi_ok;
}
bool GetI32(napi_env env, napi_value obj, const char *key, int32_t *out, int32_t dflt = 0) {
  napi_value v;
  bool has = false;
  *out = dflt;
  if (napi_has_named_property(env, obj, key, &has) != napi_ok || !has) return true;
  if (napi_get_named_property(env, obj, key, &v) != napi_ok) return false;
  return napi_get_value_int32(env, v, out) == napi_ok;
}
bool GetBytes(napi_env env, napi_value v, uint8_t **data, size_t *len) {   // Uint8Array / Buffer
  bool is_ta = false;
  if (napi_is_typedarray(env, v, &is_ta) != napi_ok || !is_ta) return false;
  napi_typedarray_type type;
  napi_value ab;
  size_t off;
  void *p;
  if (napi_get_typedarray_info(env, v, &type, len, &p, &ab, &off) != napi_ok || type != napi_uint8_array) return false;
  *data = static_cast<uint8_t *>(p);
  return true;
}

// createContext(device: number) -> External
napi_value CreateContext(napi_env env, napi_callback_info info) {
  size_t argc = 1;
  napi_value argv[1];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr));
  int32_t device = 0;
  if (argc >= 1) napi_get_value_int32(env, argv[0], &device);
  auto *c = new Ctx();
  const int rc = stitch_ctx_create(device, &c->ctx);
  if (rc != STITCH_OK) {
    delete c;
    napi_throw_error(env, nullptr, rc == STITCH_E_NODEVICE ? "No CUDA device available: libstitch_b200 has no CPU fallback"
                                                           : "stitch_ctx_create failed");
    return nullptr;
  }
  napi_value ext;
  NAPI_OK(napi_create_external(env, c, FinalizeCtx, nullptr, &ext));
  return ext;
}

// createJob(ctx, {canvasW, canvasH, outBitDepth, outKind, alphaBlend, jpegQuality, background: Uint8Array(8), bandY0, bandY1}) -> External
napi_value CreateJob(napi_env env, napi_callback_info info) {
  size_t argc = 2;
  napi_value argv[2];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr));
  Ctx *c = nullptr;
  NAPI_OK(napi_get_value_external(env, argv[0], reinterpret_cast<void **>(&c)));
  stitch_job_desc d;
  std::memset(&d, 0, sizeof d);
  GetU32(env, argv[1], "canvasW", &d.canvas_w);
  GetU32(env, argv[1], "canvasH", &d.canvas_h);
  GetU32(env, argv[1], "outBitDepth", &d.out_bit_depth, 8);
  GetU32(env, argv[1], "outKind", &d.out_kind);
  GetU32(env, argv[1], "alphaBlend", &d.alpha_blend);
  GetU32(env, argv[1], "jpegQuality", &d.jpeg_quality, 85);   // image-concat-core.ts:1054
  GetU32(env, argv[1], "bandY0", &d.band_y0);
  GetU32(env, argv[1], "bandY1", &d.band_y1);
  napi_value bg;
  bool has = false;
  if (napi_has_named_property(env, argv[1], "background", &has) == napi_ok && has &&
      napi_get_named_property(env, argv[1], "background", &bg) == napi_ok) {
    uint8_t *p;
    size_t n;
    if (GetBytes(env, bg, &p, &n)) std::memcpy(d.background, p, n < 8 ? n : 8);
  }
  auto *j = new Job();
  j->owner = c;
  const int rc = stitch_job_create(c->ctx, &d, &j->job);
  if (rc != STITCH_OK) { delete j; return ThrowStitch(env, c, rc); }
  napi_value ext;
  NAPI_OK(napi_create_external(env, j, FinalizeJob, nullptr, &ext));
  return ext;
}

// addTile(job, {width, height, bitDepth, colorType, interlace, filtered, dstX, dstY, srcOffX, srcOffY, visW, visH, zIndex,
//               imageIdx}, data: Uint8Array [, plte: Uint8Array, trns?: Uint8Array])
// Synchronous: a JS ArrayBuffer is pageable memory, so the bytes are staged before the call returns.
napi_value AddTile(napi_env env, napi_callback_info info) {
  size_t argc = 5;
  napi_value argv[5];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr));
  Job *j = nullptr;
  NAPI_OK(napi_get_value_external(env, argv[0], reinterpret_cast<void **>(&j)));
  stitch_tile_desc t;
  std::memset(&t, 0, sizeof t);
  GetU32(env, argv[1], "width", &t.width);
  GetU32(env, argv[1], "height", &t.height);
  GetU32(env, argv[1], "bitDepth", &t.bit_depth, 8);
  GetU32(env, argv[1], "colorType", &t.color_type, 6);
  GetU32(env, argv[1], "interlace", &t.interlace);
  GetU32(env, argv[1], "filtered", &t.filtered, 1);
  GetI32(env, argv[1], "dstX", &t.dst_x);
  GetI32(env, argv[1], "dstY", &t.dst_y);
  GetU32(env, argv[1], "srcOffX", &t.src_off_x);
  GetU32(env, argv[1], "srcOffY", &t.src_off_y);
  GetU32(env, argv[1], "visW", &t.vis_w, t.width);
  GetU32(env, argv[1], "visH", &t.vis_h, t.height);
  GetI32(env, argv[1], "zIndex", &t.z_index);
  GetU32(env, argv[1], "imageIdx", &t.image_idx);
  uint8_t *data = nullptr, *plte = nullptr, *trns = nullptr;
  size_t len = 0, plte_len = 0, trns_len = 0;
  if (!GetBytes(env, argv[2], &data, &len)) { napi_throw_type_error(env, nullptr, "addTile: data must be a Uint8Array"); return nullptr; }
  int rc;
  if (argc >= 4 && GetBytes(env, argv[3], &plte, &plte_len)) {   // extension: colour type 3 with its palette
    if (argc >= 5) GetBytes(env, argv[4], &trns, &trns_len);
    rc = stitch_job_add_tile_palette(j->job, &t, data, len, plte, plte_len, trns, trns_len);
  } else {
    rc = stitch_job_add_tile(j->job, &t, data, len);
  }
  if (rc != STITCH_OK) return ThrowStitch(env, j->owner, rc);
  return nullptr;
}

// addTileJpeg(job, placement, file: Uint8Array): a baseline JPEG file as a tile, decoded on the device (SURVEY 8f.4;
// replaces decodeJpeg + the row hand-over, src/decoders/jpeg-decoder.ts:239-281, :317-347).  Throws "Invalid JPEG: ..."
// or "Unsupported JPEG: ..." (the caller then falls back to the reference's own decoders and addTile).
napi_value AddTileJpeg(napi_env env, napi_callback_info info) {
  size_t argc = 3;
  napi_value argv[3];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr));
  Job *j = nullptr;
  NAPI_OK(napi_get_value_external(env, argv[0], reinterpret_cast<void **>(&j)));
  uint8_t *data = nullptr;
  size_t len = 0;
  if (!GetBytes(env, argv[2], &data, &len)) { napi_throw_type_error(env, nullptr, "addTileJpeg: file must be a Uint8Array"); return nullptr; }
  stitch_tile_desc t;
  std::memset(&t, 0, sizeof t);
  uint32_t w = 0, h = 0, ch = 0;
  int rc = stitch_jpeg_info(j->owner->ctx, data, len, &w, &h, &ch);
  if (rc != STITCH_OK) return ThrowStitch(env, j->owner, rc);
  GetI32(env, argv[1], "dstX", &t.dst_x);
  GetI32(env, argv[1], "dstY", &t.dst_y);
  GetU32(env, argv[1], "srcOffX", &t.src_off_x);
  GetU32(env, argv[1], "srcOffY", &t.src_off_y);
  GetU32(env, argv[1], "visW", &t.vis_w, w);
  GetU32(env, argv[1], "visH", &t.vis_h, h);
  GetI32(env, argv[1], "zIndex", &t.z_index);
  GetU32(env, argv[1], "imageIdx", &t.image_idx);
  rc = stitch_job_add_tile_jpeg(j->job, &t, data, len);
  if (rc != STITCH_OK) return ThrowStitch(env, j->owner, rc);
  return nullptr;
}

// jpegInfo(ctx, file) -> { width, height, channels }: parseJpegHeader + getHeader, src/decoders/jpeg-decoder.ts:46-94, :298-315
napi_value JpegInfo(napi_env env, napi_callback_info info) {
  size_t argc = 2;
  napi_value argv[2];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr));
  Ctx *c = nullptr;
  NAPI_OK(napi_get_value_external(env, argv[0], reinterpret_cast<void **>(&c)));
  uint8_t *data = nullptr;
  size_t len = 0;
  if (!GetBytes(env, argv[1], &data, &len)) { napi_throw_type_error(env, nullptr, "jpegInfo: file must be a Uint8Array"); return nullptr; }
  uint32_t w = 0, h = 0, ch = 0;
  const int rc = stitch_jpeg_info(c->ctx, data, len, &w, &h, &ch);
  if (rc != STITCH_OK) return ThrowStitch(env, c, rc);
  napi_value obj, v;
  NAPI_OK(napi_create_object(env, &obj));
  NAPI_OK(napi_create_uint32(env, w, &v)); NAPI_OK(napi_set_named_property(env, obj, "width", v));
  NAPI_OK(napi_create_uint32(env, h, &v)); NAPI_OK(napi_set_named_property(env, obj, "height", v));
  NAPI_OK(napi_create_uint32(env, ch, &v)); NAPI_OK(napi_set_named_property(env, obj, "channels", v));
  return obj;
}

// runBand(job, y0, y1, out: Uint8Array) -> Promise<number of bytes written>
// The band is rendered on a libuv worker thread: the event loop is never blocked (calls on one job are
// serialised by the host, which awaits bands in ascending y0 before feeding deflate).
struct BandWork {
  napi_async_work work = nullptr;
  napi_deferred deferred = nullptr;
  napi_ref out_ref = nullptr;   // keeps the output buffer alive while the device writes into it
  Job *job = nullptr;
  uint32_t y0 = 0, y1 = 0;
  uint8_t *out = nullptr;
  size_t cap = 0, len = 0;
  int rc = 0;
  std::string error;
};
void RunBandExecute(napi_env, void *data) {
  auto *w = static_cast<BandWork *>(data);
  w->rc = stitch_job_run_band(w->job->job, w->y0, w->y1, w->out, w->cap, &w->len, nullptr);
  if (w->rc != STITCH_OK) {   // read the message on this thread, right after the failing call
    const char *msg = stitch_last_error(w->job->owner->ctx);
    w->error = msg && *msg ? msg : ("libstitch_b200 error " + std::to_string(w->rc));
  }
}
void RunBandComplete(napi_env env, napi_status, void *data) {
  auto *w = static_cast<BandWork *>(data);
  if (w->rc == STITCH_OK) {
    napi_value n;
    napi_create_double(env, static_cast<double>(w->len), &n);
    napi_resolve_deferred(env, w->deferred, n);
  } else {
    napi_value msg, err;
    napi_create_string_utf8(env, w->error.c_str(), NAPI_AUTO_LENGTH, &msg);
    napi_create_error(env, nullptr, msg, &err);
    napi_reject_deferred(env, w->deferred, err);
  }
  napi_delete_reference(env, w->out_ref);
  napi_delete_async_work(env, w->work);
  delete w;
}
napi_value RunBand(napi_env env, napi_callback_info info) {
  size_t argc = 4;
  napi_value argv[4];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr));
  auto *w = new BandWork();
  NAPI_OK(napi_get_value_external(env, argv[0], reinterpret_cast<void **>(&w->job)));
  napi_get_value_uint32(env, argv[1], &w->y0);
  napi_get_value_uint32(env, argv[2], &w->y1);
  if (!GetBytes(env, argv[3], &w->out, &w->cap)) { delete w; napi_throw_type_error(env, nullptr, "runBand: out must be a Uint8Array"); return nullptr; }
  napi_value promise, name;
  NAPI_OK(napi_create_reference(env, argv[3], 1, &w->out_ref));
  NAPI_OK(napi_create_promise(env, &w->deferred, &promise));
  NAPI_OK(napi_create_string_utf8(env, "stitch_b200.runBand", NAPI_AUTO_LENGTH, &name));
  NAPI_OK(napi_create_async_work(env, nullptr, name, RunBandExecute, RunBandComplete, w, &w->work));
  NAPI_OK(napi_queue_async_work(env, w->work));
  return promise;
}

// jpegHeader(width, height, quality) -> Uint8Array (SOI .. SOS: what the reference's encoder emits with its first strip)
napi_value JpegHeader(napi_env env, napi_callback_info info) {
  size_t argc = 3;
  napi_value argv[3];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr));
  uint32_t w = 0, h = 0, q = 85;
  napi_get_value_uint32(env, argv[0], &w); napi_get_value_uint32(env, argv[1], &h); napi_get_value_uint32(env, argv[2], &q);
  uint8_t tmp[2048];
  size_t n = 0;
  if (stitch_jpeg_header(w, h, q, tmp, sizeof tmp, &n) != STITCH_OK) { napi_throw_range_error(env, nullptr, "Invalid JPEG dimensions or quality"); return nullptr; }
  void *p;
  napi_value ab, ta;
  NAPI_OK(napi_create_arraybuffer(env, n, &p, &ab));
  std::memcpy(p, tmp, n);
  NAPI_OK(napi_create_typedarray(env, napi_uint8_array, n, ab, 0, &ta));
  return ta;
}

// destroyJob(job), destroyContext(ctx): eager release (the finalisers also do it when the externals are collected)
napi_value DestroyJob(napi_env env, napi_callback_info info) {
  size_t argc = 1;
  napi_value argv[1];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr));
  Job *j = nullptr;
  NAPI_OK(napi_get_value_external(env, argv[0], reinterpret_cast<void **>(&j)));
  if (j->job) { stitch_job_destroy(j->job); j->job = nullptr; }
  return nullptr;
}
napi_value DestroyContext(napi_env env, napi_callback_info info) {
  size_t argc = 1;
  napi_value argv[1];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr));
  Ctx *c = nullptr;
  NAPI_OK(napi_get_value_external(env, argv[0], reinterpret_cast<void **>(&c)));
  if (c->ctx) { stitch_ctx_destroy(c->ctx); c->ctx = nullptr; }
  return nullptr;
}

napi_value Init(napi_env env, napi_value exports) {
  const napi_property_descriptor props[] = {
      {"createContext", nullptr, CreateContext, nullptr, nullptr, nullptr, napi_default, nullptr},
      {"createJob", nullptr, CreateJob, nullptr, nullptr, nullptr, napi_default, nullptr},
      {"addTile", nullptr, AddTile, nullptr, nullptr, nullptr, napi_default, nullptr},
      {"addTileJpeg", nullptr, AddTileJpeg, nullptr, nullptr, nullptr, napi_default, nullptr},
      {"jpegInfo", nullptr, JpegInfo, nullptr, nullptr, nullptr, napi_default, nullptr},
      {"runBand", nullptr, RunBand, nullptr, nullptr, nullptr, napi_default, nullptr},
      {"jpegHeader", nullptr, JpegHeader, nullptr, nullptr, nullptr, napi_default, nullptr},
      {"destroyJob", nullptr, DestroyJob, nullptr, nullptr, nullptr, napi_default, nullptr},
      {"destroyContext", nullptr, DestroyContext, nullptr, nullptr, nullptr, napi_default, nullptr},
  };
  napi_define_properties(env, exports, sizeof props / sizeof props[0], props);
  return exports;
}

}  // namespace

NAPI_MODULE(NODE_GYP_MODULE_NAME, Init)
