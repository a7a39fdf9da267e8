// vitsb_flow_reverse_noalloc: the reverse flow (fused coupling-layer plan) for callers that cannot allocate.
//
// No counterpart in the reference.  Some callers cannot call cudaMalloc
// (code under CUDA-graph capture, tight serving loops, harnesses that forbid it) and therefore cannot provide the
// workspace vitsb_flow_reverse wants.  This variant keeps the same kernels' scratch - the time-major fp32 flow state,
// its [hi | lo] operand copy, the flow conditioning vectors and the device copy of the lengths - in ONE file-scope
// device array of VITSB_NOALLOC_SCRATCH_BYTES (8 MiB), which every GPU that loads the library holds until exit.
//
//   * The call path below contains no cudaMalloc / cudaFree / stream or graph creation: only cudaSetDevice,
//     cudaGetSymbolAddress (once per device), cudaMemcpyAsync of the B lengths and kernel launches.
//   * In host code `g_noalloc_scratch` is a host-side stand-in, NOT a device address; the address the kernels get comes
//     from cudaGetSymbolAddress, taken once per device with that device current.
//   * Every call shares the array and it keeps the last call's contents.  Nothing in it has to start at zero: each
//     region a kernel reads (len, cond, zf, zop for rows [0, T)) is fully written earlier in the same call, rows beyond
//     len[b] included (the pack kernel stores zeros there).  Calls that overlap on the GPU (different streams, or
//     different handles on one device) race on the array: serialise them.
#include <cstring>
#include <mutex>

#include "engine.h"

namespace vitsb {

constexpr size_t kNoAllocScratchBytes = (size_t)VITSB_NOALLOC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

namespace {

std::mutex g_addr_mu;
void* g_addr[64] = {};          // device address of g_noalloc_scratch, per device ordinal

// device address of the scratch array on `device` (made current here); never the symbol itself
cudaError_t scratch_address(int device, void** out) {
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return e;
  if (device < 0 || device >= 64) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lk(g_addr_mu);
  if (!g_addr[device]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    g_addr[device] = p;
  }
  *out = g_addr[device];
  return cudaSuccess;
}

size_t up256(size_t x) { return (x + 255) / 256 * 256; }

}  // namespace

// bytes of the static scratch a (B, T) call needs; 0 when the handle's plan is not the fused-coupling one
size_t noalloc_flow_bytes(const Engine& e, int B, int T) {
  if (!e.flow_fused) return 0;
  const size_t BT = (size_t)B * T, IC = e.cfg.inter_channels;
  const size_t ncond = (size_t)e.cfg.n_flows * e.cfg.flow_n_layers * 2 * e.cfg.hidden_channels;
  return up256(sizeof(int) * (size_t)B) + up256(4 * (size_t)B * ncond) + up256(4 * BT * IC) + up256(2 * BT * IC * 2);
}

int noalloc_flow_reverse(Engine& e, const float* z_p, const float* g, const int32_t* len, int B, int T, float* z,
                         cudaStream_t stream) {
  if (!e.loaded) return e.fail(VITSB_ERR_WEIGHTS, "weights not loaded");
  if (!z_p || !g || !z) return e.fail(VITSB_ERR_BAD_ARG, "null pointer");
  if (B < 1 || T < 1 || B > kMaxBatchRows) return e.fail(VITSB_ERR_BAD_ARG, "B must be in [1, 512] and T_max >= 1");
  if (!e.flow_fused)
    return e.fail(VITSB_ERR_UNSUPPORTED, "vitsb_flow_reverse_noalloc needs the fused coupling-layer plan (fp16 / bf16 precision, "
                                         "base flow shape); use vitsb_flow_reverse with a caller workspace");
  if (len) for (int b = 0; b < B; ++b)
    if (len[b] < 0 || len[b] > T) return e.fail(VITSB_ERR_BAD_ARG, "len[b] out of [0, T_max]");
  const size_t need = noalloc_flow_bytes(e, B, T);
  if (need > kNoAllocScratchBytes)
    return e.fail(VITSB_ERR_WORKSPACE, "vitsb_flow_reverse_noalloc: B x T_max needs " + std::to_string(need) + " bytes of scratch, the static "
                                       "array holds " + std::to_string(kNoAllocScratchBytes) + " (VITSB_NOALLOC_SCRATCH_BYTES); use "
                                       "vitsb_flow_reverse with a caller workspace");
  void* base_v = nullptr;
  cudaError_t ce = scratch_address(e.device, &base_v);
  if (ce != cudaSuccess) return e.fail(VITSB_ERR_CUDA, std::string("noalloc scratch address: ") + cudaGetErrorString(ce));

  const int IC = e.cfg.inter_channels, nl = e.cfg.flow_n_layers;
  const int ncond = e.cfg.n_flows * nl * 2 * e.cfg.hidden_channels;      // the flow's conditioning columns come first in the packed matrix
  const size_t BT = (size_t)B * T;
  char* base = reinterpret_cast<char*>(base_v);
  size_t off = 0;
  int* d_len = reinterpret_cast<int*>(base + off);   off += up256(sizeof(int) * (size_t)B);
  float* d_cond = reinterpret_cast<float*>(base + off); off += up256(4 * (size_t)B * ncond);
  float* zf = reinterpret_cast<float*>(base + off);  off += up256(4 * BT * IC);
  void* zop = base + off;

  if (len) ce = cudaMemcpyAsync(d_len, len, sizeof(int) * (size_t)B, cudaMemcpyHostToDevice, stream);
  const int* dl = len ? d_len : nullptr;
  const LaunchOpts lo = e.lopts();
  if (ce == cudaSuccess) ce = launch_pack_time_major(z_p, B, IC, T, dl, zf, zop, e.op, 1, stream);
  if (ce == cudaSuccess) ce = launch_cond_gemv(e.d_condW, e.d_condB, g, B, ncond, e.cfg.gin_channels, d_cond, stream);
  const int t_out = flow_fused_tile_rows(nl);
  int tiles = 0;
  for (int b = 0; b < B; ++b) {
    const int rows = len ? len[b] : T;
    tiles += (rows + t_out - 1) / t_out;
  }
  for (size_t i = 0; i < e.couplings.size() && ce == cudaSuccess; ++i) {
    FlowFusedArgs fa{};
    fa.zf = zf; fa.zop = zop; fa.w = e.d_w; fa.bias = e.d_bias;
    fa.cond = d_cond; fa.cond_bstride = ncond;
    fa.fw = e.couplings[i].ff; fa.n_layers = nl;
    fa.T = T; fa.B = B; fa.len = dl; fa.m_tiles = tiles;
    if (fa.fw.cond_off + nl * 2 * e.cfg.hidden_channels > ncond) return e.fail(VITSB_ERR_UNSUPPORTED, "conditioning layout");
    ce = launch_flow_fused(fa, e.op, lo, stream);
  }
  if (ce == cudaSuccess) ce = launch_unpack_channel_major(zf, B, IC, T, z, stream);
  if (ce != cudaSuccess) return e.fail(VITSB_ERR_CUDA, std::string("vitsb_flow_reverse_noalloc: ") + cudaGetErrorString(ce));
  return VITSB_OK;
}

}  // namespace vitsb

extern "C" {

size_t vitsb_noalloc_scratch_bytes(void) { return vitsb::kNoAllocScratchBytes; }

int vitsb_flow_reverse_noalloc(vitsb_handle* h, const float* z_p, const float* g, const int32_t* len, int B, int T_max,
                               float* z, vitsb_stream stream) {
  if (!h) return VITSB_ERR_BAD_ARG;
  std::lock_guard<std::mutex> lk(h->eng.mu);
  return vitsb::noalloc_flow_reverse(h->eng, z_p, g, len, B, T_max, z, (cudaStream_t)stream);
}

}  // extern "C"
