// Host-side support for libqi_b200.so: driver entry point for tensor-map encoding, device query,
// and the integer (bit-exact) host tables of the path: MSRoPE position ids and the pack permutation.
#include "qi_common.cuh"
#include "../../include/qi_b200.h"
#include <cudaTypedefs.h>
#include <cstdlib>
#include <mutex>
#include <string>

static std::string g_last_error;

static void set_err(const std::string& s) { g_last_error = s; }
// for the other translation units of the library
void qi_set_last_error(const char* msg) { g_last_error = msg ? msg : ""; }

// Keys the persisted GEMM schedules (ops._persist_key): bump it with every change to the GEMM kernels or their
// scheduling, then re-tune and re-pin vllm_omni_draft_b200/tuned_schedules.json (tools/pin_schedules.py;
// tests/test_cpu.py::test_shipped_schedules_are_pinned_to_this_gemm_source fails until that is done).
extern "C" int qi_version(void) { return 101; }
extern "C" const char* qi_last_cuda_error(void) { return g_last_error.c_str(); }

// Per-device caches (a process may drive several GPUs; the current device decides).  Guarded by one mutex: the
// library's other process-wide knobs (qi_gemm_set_tile_n / _cta_pair / _schedule, qi_attention_set_variant) are
// documented as single-threaded setup calls, but these two sit on every launch path.
static std::mutex g_dev_mutex;
static int g_sms[64];
static bool g_sms_init = false;

int qi_num_sms() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
    set_err(std::string("no CUDA device: ") + cudaGetErrorString(cudaGetLastError()));
    return -1;
  }
  std::lock_guard<std::mutex> lock(g_dev_mutex);
  if (!g_sms_init) {
    for (int i = 0; i < 64; ++i) g_sms[i] = -2;
    g_sms_init = true;
  }
  if (g_sms[dev] == -2) {
    int n = 0, major = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) {
      set_err(std::string("no CUDA device: ") + cudaGetErrorString(cudaGetLastError()));
      return -1;  // do not cache: a later call may run after a device appears
    }
    cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    if (major != 10) {
      set_err("libqi_b200 requires an sm_100 (B200) device; there is no fallback path");
      return -1;
    }
    g_sms[dev] = n;
  }
  return g_sms[dev];
}

int qi_set_max_smem(const void* kernel, int bytes) {
  struct Entry { const void* k; unsigned long long devmask; };
  static Entry table[256];
  static int n = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return QI_ERR_CUDA;
  std::lock_guard<std::mutex> lock(g_dev_mutex);
  Entry* e = nullptr;
  for (int i = 0; i < n; ++i)
    if (table[i].k == kernel) e = &table[i];
  if (e && (e->devmask >> dev) & 1ull) return QI_OK;
  if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess) {
    set_err(std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(cudaGetLastError()));
    return QI_ERR_CUDA;
  }
  if (!e && n < 256) {
    table[n].k = kernel;
    table[n].devmask = 0;
    e = &table[n++];
  }
  if (e) e->devmask |= 1ull << dev;
  return QI_OK;
}
extern "C" int qi_device_sm_count(void) { return qi_num_sms(); }

// Programmatic dependent launch switch: on by default, QI_PDL=0 in the environment or
// qi_set_pdl(0) turns the launch attribute off (the device-side griddepcontrol instructions are
// no-ops then).
static int g_pdl = -1;
int qi_pdl_enabled() {
  if (g_pdl < 0) {
    const char* e = getenv("QI_PDL");
    g_pdl = (e && e[0] == '0') ? 0 : 1;
  }
  return g_pdl;
}
extern "C" int qi_set_pdl(int on) {
  g_pdl = on ? 1 : 0;
  return QI_OK;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

int qi_encode_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                        const uint64_t* strides_bytes, const uint32_t* box) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_err("cuTensorMapEncodeTiled driver entry point unavailable");
    return QI_ERR_DRIVER;
  }
  cuuint64_t gdim[5];
  cuuint64_t gstr[4];
  cuuint32_t bdim[5];
  cuuint32_t estr[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bdim[i] = box[i];
    estr[i] = 1;
  }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i];
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base),
                  gdim, gstr, bdim, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_err("cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
    return QI_ERR_DRIVER;
  }
  return QI_OK;
}

// ---------------------------------------------------------------------------------------------
// MSRoPE integer positions (QwenEmbedRope with scale_rope=True).
// ---------------------------------------------------------------------------------------------
extern "C" int qi_rope_position_ids(const int* shapes, int num_images, int txt_len,
                                    int32_t* img_pos, int32_t* txt_pos) {
  if (!shapes || num_images <= 0 || txt_len < 0 || !img_pos || (txt_len > 0 && !txt_pos))
    return QI_ERR_ARG;
  int max_vid = 0;
  size_t o = 0;
  for (int i = 0; i < num_images; ++i) {
    const int F = shapes[3 * i], H = shapes[3 * i + 1], W = shapes[3 * i + 2];
    if (F <= 0 || H <= 0 || W <= 0) return QI_ERR_ARG;
    const int hneg = H - H / 2, wneg = W - W / 2;
    if (i + F > 4096 || hneg > 4096 || wneg > 4096) return QI_ERR_ARG;
    for (int f = 0; f < F; ++f)
      for (int h = 0; h < H; ++h)
        for (int w = 0; w < W; ++w) {
          img_pos[o++] = i + f;
          img_pos[o++] = h < hneg ? h - hneg : h - hneg;   // [-hneg..-1, 0..H/2-1]
          img_pos[o++] = w < wneg ? w - wneg : w - wneg;
        }
    const int m = H / 2 > W / 2 ? H / 2 : W / 2;
    if (m > max_vid) max_vid = m;
  }
  if (max_vid + txt_len > 4096) return QI_ERR_ARG;
  for (int t = 0; t < txt_len; ++t) {
    txt_pos[3 * t] = txt_pos[3 * t + 1] = txt_pos[3 * t + 2] = max_vid + t;
  }
  return QI_OK;
}

// _pack_latents: view(B,C,Hl/2,2,Wl/2,2).permute(0,2,4,1,3,5).reshape(B,(Hl/2)(Wl/2),C*4)
extern "C" int qi_pack_latents_index(int channels, int hl, int wl, int64_t* idx) {
  if (!idx || channels <= 0 || hl <= 0 || wl <= 0 || (hl & 1) || (wl & 1)) return QI_ERR_ARG;
  const int H2 = hl / 2, W2 = wl / 2;
  size_t o = 0;
  for (int i = 0; i < H2; ++i)
    for (int j = 0; j < W2; ++j)
      for (int c = 0; c < channels; ++c)
        for (int di = 0; di < 2; ++di)
          for (int dj = 0; dj < 2; ++dj)
            idx[o++] = ((int64_t)c * hl + (2 * i + di)) * wl + (2 * j + dj);
  return QI_OK;
}
