// extern "C" surface declared in include/vitsb200.h.
#include <cstdio>
#include <cstring>
#include <random>

#include "engine.h"

using namespace vitsb;


static thread_local std::string g_create_err;

extern "C" {

int vitsb_create(const vitsb_config* cfg, int device, vitsb_handle** out) {
  if (!cfg || !out) { g_create_err = "null argument"; return VITSB_ERR_BAD_ARG; }
  *out = nullptr;
  vitsb_handle* h = new (std::nothrow) vitsb_handle();
  if (!h) { g_create_err = "out of memory"; return VITSB_ERR_BAD_ARG; }
  h->eng.cfg = *cfg;
  h->eng.device = device;
  int rc = h->eng.validate_config();
  if (rc != VITSB_OK) { g_create_err = h->eng.err; delete h; return rc; }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || device < 0 || device >= ndev) {
    g_create_err = e != cudaSuccess ? std::string("no CUDA device: ") + cudaGetErrorString(e) : "device index out of range";
    delete h;
    return VITSB_ERR_CUDA;      // no CPU fallback, by design
  }
  *out = h;
  return VITSB_OK;
}

void vitsb_destroy(vitsb_handle* h) { delete h; }

const char* vitsb_last_error(const vitsb_handle* h) { return h ? h->eng.err.c_str() : g_create_err.c_str(); }

int vitsb_set_option(vitsb_handle* h, const char* key, int value) {
  if (!h) return VITSB_ERR_BAD_ARG;
  std::lock_guard<std::mutex> lk(h->eng.mu);
  return h->eng.set_option(key, value);
}

int vitsb_get_option(const vitsb_handle* h, const char* key, int* value) {
  if (!h) return VITSB_ERR_BAD_ARG;
  return h->eng.get_option(key, value);
}

int vitsb_load_weights(vitsb_handle* h, const vitsb_tensor_desc* tensors, int n) {
  if (!h) return VITSB_ERR_BAD_ARG;
  std::lock_guard<std::mutex> lk(h->eng.mu);
  return h->eng.load(tensors, n);
}

size_t vitsb_workspace_bytes(const vitsb_handle* h, int B, int T_max) {
  if (!h || B < 1 || T_max < 1) return 0;
  return h->eng.workspace_bytes(B, T_max);
}

static Call make_call(int what, const float* z_in, const float* g, const int32_t* len, int B, int T, int ragged, void* ws,
                      size_t ws_bytes, vitsb_stream stream) {
  Call c;
  c.what = what; c.z_in = z_in; c.g = g; c.len = len; c.B = B; c.T = T; c.ragged = ragged;
  c.ws = ws; c.ws_bytes = ws_bytes; c.stream = (cudaStream_t)stream;
  return c;
}

int vitsb_flow_reverse(vitsb_handle* h, const float* z_p, const float* g, const int32_t* len, int B, int T_max,
                       float* z, void* ws, size_t ws_bytes, vitsb_stream stream) {
  if (!h) return VITSB_ERR_BAD_ARG;
  std::lock_guard<std::mutex> lk(h->eng.mu);
  if (!z) return h->eng.fail(VITSB_ERR_BAD_ARG, "null z");
  Call c = make_call(Engine::RUN_FLOW, z_p, g, len, B, T_max, 1, ws, ws_bytes, stream);
  c.z_out = z;
  return h->eng.run(c, false, nullptr);
}

int vitsb_decode(vitsb_handle* h, const float* z, const float* g, const int32_t* len, int B, int T_max, int ragged,
                 float* wav, void* ws, size_t ws_bytes, vitsb_stream stream) {
  if (!h) return VITSB_ERR_BAD_ARG;
  std::lock_guard<std::mutex> lk(h->eng.mu);
  Call c = make_call(Engine::RUN_DEC, z, g, len, B, T_max, ragged, ws, ws_bytes, stream);
  c.wav = wav;
  return h->eng.run(c, false, nullptr);
}

int vitsb_synthesize(vitsb_handle* h, const float* z_p, const float* g, const int32_t* len, int B, int T_max, int ragged,
                     float* z_out, float* wav, void* ws, size_t ws_bytes, vitsb_stream stream) {
  if (!h) return VITSB_ERR_BAD_ARG;
  std::lock_guard<std::mutex> lk(h->eng.mu);
  Call c = make_call(Engine::RUN_FLOW | Engine::RUN_DEC, z_p, g, len, B, T_max, ragged, ws, ws_bytes, stream);
  c.z_out = z_out; c.wav = wav;
  return h->eng.run(c, false, nullptr);
}

int vitsb_synthesize_pcm16(vitsb_handle* h, const float* z_p, const float* g, const int32_t* len, int B, int T_max, int ragged,
                           float volume, int16_t* pcm, void* ws, size_t ws_bytes, vitsb_stream stream) {
  if (!h) return VITSB_ERR_BAD_ARG;
  std::lock_guard<std::mutex> lk(h->eng.mu);
  if (!pcm) return h->eng.fail(VITSB_ERR_BAD_ARG, "null pcm");
  Call c = make_call(Engine::RUN_FLOW | Engine::RUN_DEC, z_p, g, len, B, T_max, ragged, ws, ws_bytes, stream);
  c.pcm = reinterpret_cast<short*>(pcm); c.volume = volume;      // call-local: nothing about the output form lives in the handle
  return h->eng.run(c, false, nullptr);
}

int vitsb_synthesize_durations(vitsb_handle* h, const float* m_p, const float* s_p, const float* logw, const float* g,
                               const int32_t* x_len, const float* noise, long long noise_stride, int B, int Tx_max, int T_cap,
                               float length_scale, int ragged, float* z_out, float* wav, int16_t* pcm, float volume,
                               int32_t* y_len, void* ws, size_t ws_bytes, vitsb_stream stream) {
  if (!h) return VITSB_ERR_BAD_ARG;
  std::lock_guard<std::mutex> lk(h->eng.mu);
  if (!wav && !pcm) return h->eng.fail(VITSB_ERR_BAD_ARG, "null wav / pcm");
  DurFront d;
  d.m_p = m_p; d.s_p = s_p; d.logw = logw; d.noise = noise; d.noise_stride = noise_stride; d.x_len = x_len; d.Tx = Tx_max;
  d.scale = length_scale; d.y_len_out = reinterpret_cast<int*>(y_len);
  Call c = make_call(Engine::RUN_FLOW | Engine::RUN_DEC, nullptr, g, nullptr, B, T_cap, ragged, ws, ws_bytes, stream);
  c.dur = &d; c.z_out = z_out; c.wav = wav; c.pcm = reinterpret_cast<short*>(pcm); c.volume = volume;
  return h->eng.run(c, false, nullptr);
}

int vitsb_length_regulate(vitsb_handle* h, const float* attn, const float* m_p, const float* s_p, const float* noise,
                          int B, int T, int Tx, float* z_p, vitsb_stream stream) {
  if (!h) return VITSB_ERR_BAD_ARG;
  std::lock_guard<std::mutex> lk(h->eng.mu);
  if (!attn || !m_p || !s_p || !noise || !z_p) return h->eng.fail(VITSB_ERR_BAD_ARG, "null pointer");
  if (B < 1 || T < 1 || Tx < 1) return h->eng.fail(VITSB_ERR_BAD_ARG, "B, T and Tx must be >= 1");
  if (B > 65535) return h->eng.fail(VITSB_ERR_BAD_ARG, "B too large");
  cudaError_t e = cudaSetDevice(h->eng.device);
  if (e == cudaSuccess)
    e = vitsb::launch_length_regulate(attn, m_p, s_p, noise, B, h->eng.cfg.inter_channels, T, Tx, z_p, (cudaStream_t)stream);
  if (e != cudaSuccess) return h->eng.fail(VITSB_ERR_CUDA, std::string("length_regulate: ") + cudaGetErrorString(e));
  return VITSB_OK;
}

int vitsb_resample_poly(vitsb_handle* h, const float* x, int B, long long n_in, long long x_stride, int up, int down,
                        const double* taps, int n_taps, long long n_pre_remove, long long n_out, float* y, int16_t* pcm,
                        float volume, long long y_stride, vitsb_stream stream) {
  if (!h) return VITSB_ERR_BAD_ARG;
  std::lock_guard<std::mutex> lk(h->eng.mu);
  if (!x || !taps || (!y && !pcm)) return h->eng.fail(VITSB_ERR_BAD_ARG, "null pointer");
  if (B < 1 || B > 65535) return h->eng.fail(VITSB_ERR_BAD_ARG, "B must be in [1, 65535]");
  if (up < 1 || down < 1 || n_taps < 1 || n_in < 1 || n_out < 0 || n_pre_remove < 0)
    return h->eng.fail(VITSB_ERR_BAD_ARG, "up, down, n_taps, n_in must be >= 1 and n_out, n_pre_remove >= 0");
  if (x_stride < n_in || y_stride < n_out) return h->eng.fail(VITSB_ERR_BAD_ARG, "row strides must cover n_in / n_out");
  if (n_out > 0x7fffffffLL * 256) return h->eng.fail(VITSB_ERR_BAD_ARG, "n_out too large");
  cudaError_t e = cudaSetDevice(h->eng.device);
  if (e == cudaSuccess)
    e = vitsb::launch_resample_poly(x, B, n_in, x_stride, up, down, taps, n_taps, n_pre_remove, n_out, y, pcm, volume, y_stride,
                                    (cudaStream_t)stream);
  if (e != cudaSuccess) return h->eng.fail(VITSB_ERR_CUDA, std::string("resample_poly: ") + cudaGetErrorString(e));
  return VITSB_OK;
}

int vitsb_total_upsample(const vitsb_handle* h) { return h ? h->eng.up_total : 0; }

int vitsb_launches_per_call(const vitsb_handle* h, int B, int T_max) {
  if (!h) return 0;
  int n = 0;
  Engine* e = const_cast<Engine*>(&h->eng);
  std::lock_guard<std::mutex> lk(e->mu);
  Call c;
  c.what = Engine::RUN_FLOW | Engine::RUN_DEC; c.B = B; c.T = T_max; c.ragged = 1;
  if (e->run(c, true, &n) != VITSB_OK) return 0;
  return n;
}

double vitsb_flop_per_frame(const vitsb_handle* h) { return h ? h->eng.flop_per_frame : 0.0; }

int vitsb_profile(vitsb_handle* h, int enable) {
  if (!h) return VITSB_ERR_BAD_ARG;
  std::lock_guard<std::mutex> lk(h->eng.mu);
  h->eng.profile = enable != 0;
  h->eng.prof.clear();
  return VITSB_OK;
}

int vitsb_profile_read(const vitsb_handle* h, int32_t* tags, double* ms, int64_t* launches, double* flops, int cap) {
  if (!h) return 0;
  int n = 0;
  for (const auto& kv : h->eng.prof) {
    if (n >= cap) break;
    if (tags) tags[n] = kv.first;
    if (ms) ms[n] = kv.second.ms;
    if (launches) launches[n] = kv.second.launches;
    if (flops) flops[n] = kv.second.flops;
    ++n;
  }
  return n;
}

const char* vitsb_build_info(void) {
  return "vitsb200;arch=sm_100a;tensor=tcgen05.mma.kind::f16+TMEM;copy=TMA(cp.async.bulk.tensor);flow=fused-coupling;resblock=fused;verify=fp32-ffma;cpu_fallback=none";
}

// ---------------------------------------------------------------------------
// Bring-up self-test: one conv configuration on the tcgen05 path against the
// CUDA-core fp32 kernel on identical (operand-rounded) data.
// ---------------------------------------------------------------------------
int vitsb_selftest_conv(int device, int precision, int Cin, int N, int k, int dil, int L, int B, int gate, uint32_t seed,
                        double* max_err, double* ref_scale) {
  if (!max_err || !ref_scale) return VITSB_ERR_BAD_ARG;
  if (precision != VITSB_PREC_FP16 && precision != VITSB_PREC_BF16) return VITSB_ERR_BAD_ARG;
  const int op = precision == VITSB_PREC_BF16 ? OP_BF16 : OP_F16;
  const int NT = N <= kMaxTileN ? N : kMaxTileN;
  const bool staged_out = gate == 2;        // gate = 2: plain 16-bit output only (the staged TMA-store epilogue of the upsamplers)
  if (staged_out) gate = 0;
  if (!conv_tc_supported(Cin, Cin, N, NT, k) || (gate && NT % 32)) return VITSB_ERR_UNSUPPORTED;
  if (cudaSetDevice(device) != cudaSuccess) return VITSB_ERR_CUDA;
  std::mt19937 rng(seed);
  std::uniform_real_distribution<float> U(-1.f, 1.f);
  const size_t n_in = (size_t)B * L * Cin, n_w = (size_t)N * k * Cin, Nout = gate ? N / 2 : N, n_out = (size_t)B * L * Nout;
  std::vector<float> in(n_in), w(n_w), bias((size_t)B * N), res(n_out);
  const float ws = 1.0f / std::sqrt((float)(k * Cin));
  auto rnd = [&](float v) { return op == OP_BF16 ? __bfloat162float(__float2bfloat16_rn(v)) : __half2float(__float2half_rn(v)); };
  for (auto& x : in) x = rnd(U(rng));
  for (auto& x : w) x = rnd(U(rng) * ws * 1.7f);
  for (auto& x : bias) x = U(rng) * 0.3f;
  for (auto& x : res) x = U(rng);
  std::vector<uint16_t> in16(n_in), w16(n_w);
  for (size_t i = 0; i < n_in; ++i) {
    if (op == OP_BF16) { __nv_bfloat16 h = __float2bfloat16_rn(in[i]); std::memcpy(&in16[i], &h, 2); }
    else { __half h = __float2half_rn(in[i]); std::memcpy(&in16[i], &h, 2); }
  }
  for (size_t i = 0; i < n_w; ++i) {
    if (op == OP_BF16) { __nv_bfloat16 h = __float2bfloat16_rn(w[i]); std::memcpy(&w16[i], &h, 2); }
    else { __half h = __float2half_rn(w[i]); std::memcpy(&w16[i], &h, 2); }
  }
  std::vector<int> len(B);
  for (int b = 0; b < B; ++b) len[b] = b == 0 ? L : (int)(L * (0.35 + 0.6 * (b % 3) / 2.0));
  float *d_in32, *d_w32, *d_bias, *d_res, *d_o1, *d_o2; void *d_in16, *d_w16, *d_op1, *d_op2; int* d_len;
  cudaMalloc(&d_in32, n_in * 4); cudaMalloc(&d_w32, n_w * 4); cudaMalloc(&d_bias, bias.size() * 4); cudaMalloc(&d_res, n_out * 4);
  cudaMalloc(&d_o1, n_out * 4); cudaMalloc(&d_o2, n_out * 4); cudaMalloc(&d_in16, n_in * 2); cudaMalloc(&d_w16, n_w * 2);
  cudaMalloc(&d_op1, n_out * 4); cudaMalloc(&d_op2, n_out * 2); cudaMalloc(&d_len, B * 4);
  cudaMemcpy(d_in32, in.data(), n_in * 4, cudaMemcpyHostToDevice); cudaMemcpy(d_w32, w.data(), n_w * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(d_bias, bias.data(), bias.size() * 4, cudaMemcpyHostToDevice); cudaMemcpy(d_res, res.data(), n_out * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(d_in16, in16.data(), n_in * 2, cudaMemcpyHostToDevice); cudaMemcpy(d_w16, w16.data(), n_w * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(d_len, len.data(), B * 4, cudaMemcpyHostToDevice);
  cudaMemset(d_o1, 0xff, n_out * 4); cudaMemset(d_o2, 0xff, n_out * 4);
  ConvArgs a{};
  a.in_ld = Cin; a.in_cols = Cin; a.in_bstride = (long long)L * Cin; a.Cin = Cin; a.N = N; a.NT = NT; a.k = k; a.dil = dil; a.pad = (k - 1) / 2 * dil;
  a.L = L; a.B = B; a.len = d_len; a.len_mul = 1;
  a.epi.bias = d_bias; a.epi.bias_bstride = N; a.epi.alpha = gate ? 1.f : 0.5f; a.epi.beta = 0.75f; a.epi.out_slope = 0.1f;
  a.epi.gate = gate;
  if (!gate && !staged_out) { a.epi.res = d_res; a.epi.res_ld = (int)Nout; a.epi.res_bstride = (long long)L * Nout; }
  if (staged_out) { a.epi.bias_bstride = 0; }
  a.epi.out_f32_ld = (int)Nout; a.epi.out_f32_bstride = (long long)L * Nout;
  a.epi.out_op_ld = (int)Nout; a.epi.out_op_bstride = (long long)L * Nout;
  ConvArgs r = a;
  r.in = d_in32; r.w = d_w32; r.epi.out_f32 = d_o1; r.epi.out_op = d_op1;
  a.in = d_in16; a.w = d_w16; a.epi.out_f32 = staged_out ? nullptr : d_o2; a.epi.out_op = d_op2;
  if (staged_out) cudaMemset(d_op2, 0xff, n_out * 2);
  cudaError_t e1 = launch_conv_f32(r, 0);
  cudaError_t e2 = launch_conv_tc(a, op, LaunchOpts{}, 0);
  cudaError_t e3 = cudaDeviceSynchronize();
  int rc = VITSB_OK;
  if (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess) {
    std::fprintf(stderr, "selftest: f32=%s tc=%s sync=%s\n", cudaGetErrorString(e1), cudaGetErrorString(e2), cudaGetErrorString(e3));
    rc = VITSB_ERR_CUDA;
  } else {
    std::vector<float> o1(n_out), o2(n_out);
    cudaMemcpy(o1.data(), d_o1, n_out * 4, cudaMemcpyDeviceToHost); cudaMemcpy(o2.data(), d_o2, n_out * 4, cudaMemcpyDeviceToHost);
    if (staged_out) {
      // compare leaky_relu(reference fp32) with the 16-bit operand rows the tensor-core path stored
      std::vector<uint16_t> h(n_out);
      cudaMemcpy(h.data(), d_op2, n_out * 2, cudaMemcpyDeviceToHost);
      for (size_t i = 0; i < n_out; ++i) {
        const float v = o1[i] >= 0.f ? o1[i] : a.epi.out_slope * o1[i];
        o1[i] = rnd(v);
        if (op == OP_BF16) { __nv_bfloat16 x; std::memcpy(&x, &h[i], 2); o2[i] = __bfloat162float(x); }
        else { __half x; std::memcpy(&x, &h[i], 2); o2[i] = __half2float(x); }
      }
    }
    double me = 0, sc = 0;
    for (int b = 0; b < B; ++b) {
      // rows both paths computed: the fp32 conv works in 32- or 64-row tiles, the tensor-core conv in 128-row tiles, each up to
      // the tile that holds row len + halo - 1; rows beyond are never written by one path or the other
      for (int t = 0; t < L; ++t) {
        if ((t / 32) * 32 >= len[b] + kHaloRows) continue;
        for (size_t n = 0; n < Nout; ++n) {
          size_t i = ((size_t)b * L + t) * Nout + n;
          double d = std::fabs((double)o1[i] - (double)o2[i]);
          if (!(d <= me)) me = d;          // NaN-propagating max
          if (std::fabs(o1[i]) > sc) sc = std::fabs(o1[i]);
        }
      }
    }
    if (!(me == me)) {                      // NaN: say where (an unwritten output row, or a non-finite value)
      long long n1 = 0, n2 = 0; int fb = -1, ft = -1, fn = -1;
      for (int b = 0; b < B; ++b) for (int t = 0; t < L; ++t) {
        if ((t / 32) * 32 >= len[b] + kHaloRows) continue;
        for (size_t n = 0; n < Nout; ++n) {
          const size_t i = ((size_t)b * L + t) * Nout + n;
          if (!(o1[i] == o1[i])) ++n1;
          if (!(o2[i] == o2[i])) { if (fb < 0) { fb = b; ft = t; fn = (int)n; } ++n2; }
        }
      }
      std::fprintf(stderr, "selftest: NaN in compared rows: reference %lld, tensor-core path %lld (first at b=%d t=%d n=%d; len = %d)\n",
                   n1, n2, fb, ft, fn, fb >= 0 ? len[fb] : -1);
    }
    *max_err = me; *ref_scale = sc;
  }
  cudaFree(d_in32); cudaFree(d_w32); cudaFree(d_bias); cudaFree(d_res); cudaFree(d_o1); cudaFree(d_o2);
  cudaFree(d_in16); cudaFree(d_w16); cudaFree(d_op1); cudaFree(d_op2); cudaFree(d_len);
  return rc;
}

}  // extern "C"
