// Host-side engine: packed weights + the per-call launch plan.
#pragma once
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/vitsb200.h"
#include "kernels.h"

namespace vitsb {

// One implicit-GEMM conv after packing.
struct PackedConv {
  int Cin = 0, N = 0, NT = 0, k = 1, dil = 1, pad = 0;   // Cin = K per tap in the packed weights
  int in_cols = 0;         // physical columns of an input row (2*C for split-precision convs, else Cin)
  int in_ld = 0;           // row pitch of the input buffer in elements (0 -> in_cols); an unsplit conv over a [hi|lo] buffer reads hi only
  bool gate = false;
  int gate_group = 0;      // gated: packed columns in groups of gate_group = [tanh | sigmoid] halves (0 -> NT)
  bool split = false;      // 3-term split precision: rows [hi|lo], weights [Wh|Wh|Wl] per tap (common.h)
  size_t w_off = 0;        // element offset into the operand-typed weight pool, [N][k*Cin] K-major
  size_t bias_off = 0;     // offset into the static fp32 bias pool (packed column order)
  int cond_off = -1;       // >=0: bias = per-call conditioning vector at this column (conv bias folded in)
  double flop_per_pos = 0; // algorithmic 2*Cin*Cout*k (true taps/channels only)
  int tap_split_n = 0;     // ConvArgs::tap_split_n (upsamplers)
};

struct Coupling {          // modules.ResidualCouplingLayer (reverse, mean_only) with Flip folded in
  PackedConv pre, post;
  std::vector<PackedConv> in_layers, res, skip;
  int coff = 0;            // physical channel offset of the half updated by `post`
  FlowFusedWeights ff;     // the same layer in the fused coupling kernel's packing (flow_fused.cu); valid when Engine::flow_fused
  double flop_per_pos = 0; // algorithmic FLOP of the whole coupling per frame
};

struct ResIter { PackedConv c1, c2; };

struct Stage {             // ups[i] + its MRF (num_kernels ResBlock2 x iterations)
  PackedConv ups;
  int u = 1, C = 0;
  std::vector<std::vector<ResIter>> rb;   // [kernel][iter]
};

struct WsLayout {
  size_t len = 0, cond = 0, xlen = 0, cum = 0;      // cum: int [B][1024] cumulative durations (durations front end)
  size_t zf = 0, zop = 0, hf = 0, hop = 0, acts = 0, skipf = 0, skipop = 0;   // flow (h / acts / skip: layer-at-a-time flow only)
  size_t xin_f = 0, xb_f = 0, acc_f = 0, g_op = 0;                            // decoder, layer-at-a-time plan (fp32 build)
  size_t xin_op = 0, xb_op = 0, xb2_op = 0, br0 = 0, br1 = 0, x3a = 0, x3b = 0, nxt_op = 0;     // decoder, 16-bit streams
  size_t total = 0;
};

// One call through the C ABI (all pointers caller-owned; nothing here outlives the call).
// Device-side durations front end: the call starts from infer_p1's outputs instead of z_p.
struct DurFront {
  const float* m_p = nullptr;     // [B][inter][Tx]
  const float* s_p = nullptr;     // [B][inter][Tx]
  const float* logw = nullptr;    // [B][1][Tx]
  const float* noise = nullptr;   // [B][inter][noise_stride], or compact [inter][y_len[b]] blocks when noise_stride == 0
  long long noise_stride = 0;
  const int32_t* x_len = nullptr; // host [B] or null
  int Tx = 0;
  float scale = 1.f;              // length_scale / duration_rate
  int* y_len_out = nullptr;       // device [B], optional
};

struct Call {
  int what = 0;                   // Engine::RUN_FLOW | Engine::RUN_DEC
  const float* z_in = nullptr;    // [B][inter][T] fp32, reference layout
  const float* g = nullptr;       // [B][gin]
  const int32_t* len = nullptr;   // host [B] or null
  int B = 0, T = 0, ragged = 1;
  float* z_out = nullptr;         // optional [B][inter][T]
  float* wav = nullptr;           // [B][1][T*up] fp32 ...
  short* pcm = nullptr;           // ... or int16 PCM with the serving path's volume scaling (vits_wrap.py:195)
  float volume = 1.f;
  void* ws = nullptr; size_t ws_bytes = 0;
  cudaStream_t stream = nullptr;
  const DurFront* dur = nullptr;  // set: z_in / len are unused, T is the frame capacity, the lengths live on the device only
};

struct P1State;                 // packed text encoder + duration predictor (p1.cu)
void p1_destroy(P1State* s);
bool p1_present(const vitsb_tensor_desc* t, int n);

struct Engine {
  vitsb_config cfg{};
  P1State* p1 = nullptr;        // infer_p1 side, present when the state_dict carried enc_p.* / dp.* / emb_g.*
  int device = 0;
  int num_sms = 148;
  int op = OP_F16;
  std::string err;
  std::mutex mu;                // the C entry points serialise on it: one call at a time per handle (host side); the
                                // enqueued work is stream-ordered as usual
  bool split16 = false;         // 16-bit paths: flow + conv_pre + upsamplers run split-precision
  bool flow_in_split = false;   // WN in_layers run 3-term split precision too; default: fp16 single term, bf16 split
  bool rb_fused = false;        // 16-bit paths: ResBlock iterations run as the fused persistent kernel (rb_fused.cu)
  bool flow_fused = false;      // 16-bit paths: each coupling layer runs as one cluster kernel (flow_fused.cu)

  // options (vitsb_set_option); -1 = automatic.  Packing / plan options take effect at the next vitsb_load_weights.
  int opt_graph = 1, opt_pdl = 1, opt_tc_pair = 1, opt_rb_fused = -1, opt_flow_fused = -1, opt_flow_in_split = -1,
      opt_rb_nslab = 0, opt_ws_cap_mb = 24576, opt_ups_split_mask = -1, opt_branch_streams = 1;
  int set_option(const char* key, int value);
  int get_option(const char* key, int* value) const;
  LaunchOpts lopts() const { LaunchOpts o; o.pdl = opt_pdl != 0; o.pair = opt_tc_pair != 0; o.num_sms = num_sms; o.rb_nslab = opt_rb_nslab; return o; }

  // packed model
  bool loaded = false;
  std::vector<Coupling> couplings;       // in execution order (flows.6, .4, .2, .0)
  PackedConv conv_pre;
  std::vector<Stage> stages;
  int post_k = 7, post_C = 0;
  int Ncond = 0;
  int up_total = 1;
  double flop_per_frame = 0;

  // host staging (freed after upload)
  std::vector<float> h_w, h_bias, h_condW, h_condB, h_post;

  // device pools
  void* d_w = nullptr;          // op-typed
  float* d_bias = nullptr;
  float* d_condW = nullptr;
  float* d_condB = nullptr;
  float* d_post = nullptr;
  void free_device();

  // optional per-launch timing (cudaEvent pairs around every launch, read back after a sync)
  bool profile = false;
  struct ProfRow { double ms = 0; long long launches = 0; double flops = 0; };
  std::map<int, ProfRow> prof;     // keyed by launch tag (see engine_run.cu)

  ~Engine();
  int fail(int code, const std::string& msg) { err = msg; return code; }

  int validate_config();
  int resolve_plan();              // precision + options -> rb_fused / flow_fused / flow_in_split (at create and before every load)
  int load(const vitsb_tensor_desc* t, int n);
  WsLayout layout(int B, int T) const;
  size_t workspace_bytes(int B, int T) const;     // layout of the largest row chunk under the workspace cap
  int chunk_rows(int B, int T, size_t ws_bytes) const;

  // CUDA-graph replay of the launch plan (engine_run.cu): the ~80 kernels between the two seams of a call are
  // launch-latency bound for short batches; a call whose launch geometry (workspace, B, T, per-rate tile counts) has
  // been seen before replays an instantiated graph instead of re-issuing them.  The lengths themselves are read by
  // the kernels from device memory, so they are not part of the key.  Option "cuda_graph" = 0 disables.
  struct GraphEntry { std::vector<long long> key; int seen = 0; int launches = 0; cudaGraphExec_t exec = nullptr; unsigned long long stamp = 0; };
  std::vector<GraphEntry> graphs;
  cudaStream_t cap_stream = nullptr;
  // the MRF branches of a stage run concurrently: two side streams + fork / join events (engine_run.cu)
  cudaStream_t side_stream[2] = {nullptr, nullptr};
  cudaEvent_t ev_fork = nullptr, ev_join[2] = {nullptr, nullptr};
  bool side_failed = false;
  bool side_streams_ready();
  unsigned long long graph_clock = 0;
  void drop_graphs();
  std::vector<long long> graph_key(const Call& c) const;

  // mode bits for run()
  enum { RUN_FLOW = 1, RUN_DEC = 2 };
  // phases of one call: FRONT = length upload + seam pack + conditioning GEMV (read caller pointers),
  // MID = every conv of the flow and the decoder (workspace only), BACK = conv_post (writes the caller's waveform)
  enum { PH_FRONT = 1, PH_MID = 2, PH_BACK = 4, PH_ALL = 7 };
  int check_call(const Call& c, bool dry);
  int issue(int phases, const Call& c, bool dry, int* launches);
  int run_chunk(const Call& c, bool dry, int* launches);
  int run(const Call& c, bool dry, int* launches);
};

}  // namespace vitsb

// the opaque handle of include/vitsb200.h
struct vitsb_handle { vitsb::Engine eng; };
