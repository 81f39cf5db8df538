// Batched JPEG decode straight into device memory, for the Cluster Serving front end.
//
// The reference decodes every record on the JVM: base64 -> OpenCV imdecode -> float HWC -> CHW, one image at a time inside a
// parallel collection (zoo/.../serving/PreProcessing.scala:24-43, ClusterServing.scala:192-216).  The round-1 engine kept that
// shape (cv2.imdecode on a host thread pool) and was host-bound at ~2.5 K records/s on a GPU that classifies ten times that.
// Here a whole batch of JPEG byte strings is handed to nvJPEG's batched decoder (hardware JPEG engines when the device has
// them, GPU-assisted Huffman decoding otherwise); the decoded interleaved BGR images land in ONE device buffer at caller-given
// offsets -- exactly the packed layout `zoo_serving_preproc` (K13: resize + normalise + layout) consumes -- so pixels never
// visit host memory and Python touches a batch, not an image.
//
// libnvjpeg is resolved with dlopen at first use: libzoo_cuda.so keeps loading on a machine without it (callers then fall
// back to host decoding).
#include "../common/zoo_common.cuh"
#include <dlfcn.h>
#include <nvjpeg.h>
#include <mutex>
#include <vector>

namespace {

struct NvjApi {
  void* so = nullptr;
  nvjpegStatus_t (*CreateEx)(nvjpegBackend_t, nvjpegDevAllocator_t*, nvjpegPinnedAllocator_t*, unsigned int, nvjpegHandle_t*) = nullptr;
  nvjpegStatus_t (*Destroy)(nvjpegHandle_t) = nullptr;
  nvjpegStatus_t (*StateCreate)(nvjpegHandle_t, nvjpegJpegState_t*) = nullptr;
  nvjpegStatus_t (*StateDestroy)(nvjpegJpegState_t) = nullptr;
  nvjpegStatus_t (*GetImageInfo)(nvjpegHandle_t, const unsigned char*, size_t, int*, nvjpegChromaSubsampling_t*, int*, int*) = nullptr;
  nvjpegStatus_t (*BatchedInit)(nvjpegHandle_t, nvjpegJpegState_t, int, int, nvjpegOutputFormat_t) = nullptr;
  nvjpegStatus_t (*Batched)(nvjpegHandle_t, nvjpegJpegState_t, const unsigned char* const*, const size_t*, nvjpegImage_t*, cudaStream_t) = nullptr;
  bool ok = false;
};

NvjApi& api() {
  static NvjApi a;
  static std::once_flag once;
  std::call_once(once, [] {
    for (const char* name : {"libnvjpeg.so.12", "libnvjpeg.so", "/usr/local/cuda/lib64/libnvjpeg.so.12"}) {
      a.so = dlopen(name, RTLD_NOW | RTLD_LOCAL);
      if (a.so) break;
    }
    if (!a.so) return;
#define ZOO_NVJ(field, sym) a.field = reinterpret_cast<decltype(a.field)>(dlsym(a.so, sym)); if (!a.field) return
    ZOO_NVJ(CreateEx, "nvjpegCreateEx");
    ZOO_NVJ(Destroy, "nvjpegDestroy");
    ZOO_NVJ(StateCreate, "nvjpegJpegStateCreate");
    ZOO_NVJ(StateDestroy, "nvjpegJpegStateDestroy");
    ZOO_NVJ(GetImageInfo, "nvjpegGetImageInfo");
    ZOO_NVJ(BatchedInit, "nvjpegDecodeBatchedInitialize");
    ZOO_NVJ(Batched, "nvjpegDecodeBatched");
#undef ZOO_NVJ
    a.ok = true;
  });
  return a;
}

struct JpegBatch {
  nvjpegHandle_t handle = nullptr;
  nvjpegJpegState_t state = nullptr;
  int backend = -1;
  int batch_initialized = 0;
  int cpu_threads = 4;
  std::vector<nvjpegImage_t> outs;
};

int nvj_fail(const char* what, nvjpegStatus_t st, int line) {
  char msg[160];
  snprintf(msg, sizeof(msg), "%s failed (nvjpegStatus %d)", what, (int)st);
  zoo_set_last_error(msg, __FILE__, line);
  return ZOO_ERR_CUDA;
}

}  // namespace

// 1 when libnvjpeg could be loaded and every entry point resolved.
ZOO_API int zoo_jpeg_available() { return api().ok ? 1 : 0; }

// Create a decoder for the current device.  backend: -1 = best available (hardware engines, then GPU-assisted, then default),
// otherwise an nvjpegBackend_t value.  *out receives an opaque handle; *backend_used the backend that was created.
ZOO_API int zoo_jpeg_batch_create(int backend, int cpu_threads, void** out, int* backend_used) {
  if (!out) return ZOO_ERR_ARG;
  NvjApi& a = api();
  if (!a.ok) { zoo_set_last_error("libnvjpeg is not available on this machine", __FILE__, __LINE__); return ZOO_ERR_CUDA; }
  ZOO_CUDA_CHECK(cudaFree(0));
  JpegBatch* jb = new JpegBatch();
  jb->cpu_threads = cpu_threads > 0 ? cpu_threads : 4;
  const int order_auto[3] = {NVJPEG_BACKEND_HARDWARE, NVJPEG_BACKEND_GPU_HYBRID, NVJPEG_BACKEND_DEFAULT};
  nvjpegStatus_t st = NVJPEG_STATUS_NOT_INITIALIZED;
  for (int k = 0; k < 3; ++k) {
    const int b = backend >= 0 ? backend : order_auto[k];
    st = a.CreateEx((nvjpegBackend_t)b, nullptr, nullptr, 0, &jb->handle);
    if (st == NVJPEG_STATUS_SUCCESS) { jb->backend = b; break; }
    jb->handle = nullptr;
    if (backend >= 0) break;
  }
  if (!jb->handle) { delete jb; return nvj_fail("nvjpegCreateEx", st, __LINE__); }
  st = a.StateCreate(jb->handle, &jb->state);
  if (st != NVJPEG_STATUS_SUCCESS) { a.Destroy(jb->handle); delete jb; return nvj_fail("nvjpegJpegStateCreate", st, __LINE__); }
  if (backend_used) *backend_used = jb->backend;
  *out = jb;
  return ZOO_OK;
}

ZOO_API int zoo_jpeg_batch_destroy(void* h) {
  JpegBatch* jb = static_cast<JpegBatch*>(h);
  if (!jb) return ZOO_ERR_ARG;
  NvjApi& a = api();
  if (jb->state) a.StateDestroy(jb->state);
  if (jb->handle) a.Destroy(jb->handle);
  delete jb;
  return ZOO_OK;
}

// Width / height / channel count of every image of the batch (host call, parses headers only).
ZOO_API int zoo_jpeg_batch_info(void* h, const unsigned char* const* data, const size_t* lengths, int n, int* widths, int* heights, int* comps) {
  JpegBatch* jb = static_cast<JpegBatch*>(h);
  if (!jb || !data || !lengths || !widths || !heights || n < 0) return ZOO_ERR_ARG;
  NvjApi& a = api();
  for (int i = 0; i < n; ++i) {
    int nc = 0, ws[NVJPEG_MAX_COMPONENT], hs[NVJPEG_MAX_COMPONENT];
    nvjpegChromaSubsampling_t ss;
    nvjpegStatus_t st = a.GetImageInfo(jb->handle, data[i], lengths[i], &nc, &ss, ws, hs);
    if (st != NVJPEG_STATUS_SUCCESS) return nvj_fail("nvjpegGetImageInfo", st, __LINE__);
    widths[i] = ws[0]; heights[i] = hs[0];
    if (comps) comps[i] = nc;
  }
  return ZOO_OK;
}

// Decode n JPEG byte strings (host memory) into `out` (device memory): image i is written as interleaved 8-bit BGR
// (bgr != 0) or RGB rows of widths[i] * 3 bytes starting at out + offsets[i].  Asynchronous on `stream`.
ZOO_API int zoo_jpeg_batch_decode(void* h, const unsigned char* const* data, const size_t* lengths, int n, unsigned char* out,
                                  const long long* offsets, const int* widths, int bgr, void* stream) {
  JpegBatch* jb = static_cast<JpegBatch*>(h);
  if (!jb || !data || !lengths || !out || !offsets || !widths || n < 0) return ZOO_ERR_ARG;
  if (n == 0) return ZOO_OK;
  NvjApi& a = api();
  if (jb->batch_initialized != n) {      // the batched state is sized per batch size
    nvjpegStatus_t st = a.BatchedInit(jb->handle, jb->state, n, jb->cpu_threads, bgr ? NVJPEG_OUTPUT_BGRI : NVJPEG_OUTPUT_RGBI);
    if (st != NVJPEG_STATUS_SUCCESS) return nvj_fail("nvjpegDecodeBatchedInitialize", st, __LINE__);
    jb->batch_initialized = n;
  }
  jb->outs.resize((size_t)n);
  for (int i = 0; i < n; ++i) {
    nvjpegImage_t im{};
    im.channel[0] = out + offsets[i];
    im.pitch[0] = (size_t)widths[i] * 3;
    jb->outs[(size_t)i] = im;
  }
  nvjpegStatus_t st = a.Batched(jb->handle, jb->state, data, lengths, jb->outs.data(), (cudaStream_t)stream);
  if (st != NVJPEG_STATUS_SUCCESS) return nvj_fail("nvjpegDecodeBatched", st, __LINE__);
  return ZOO_OK;
}
