// K-P: the persistent per-sample trainer.  One cooperative launch per sample runs
// the whole convergence loop of ann_train_BP/BPM / snn_train_BP/BPM
// (src/ann.c:2281-2467, src/snn.c:1417-1595) on the device: the host is involved
// once per sample instead of ~40 times per iteration.
//
// Per iteration (weight layers 0..NL-1):
//   top-down   l = NL-1..1 : column tasks -> partial W_l^T.delta_l from the OLD
//                            weights, BP/BPM update of W_l in the same pass, partials
//                            pushed to the rank owning the matching rows of layer l-1
//                            (reduce-scatter by peer stores)               [barrier]
//   bottom-up  l = 0       : row tasks: update W_0 and forward in one pass  [barrier]
//              l = 1..NL-1 : row tasks: forward with the new weights; every result is
//                            stored into all ranks' activation buffers
//                            (all-gather by peer stores)                    [barrier]
//   every CTA then evaluates softmax / error / argmax / loop test redundantly.
// => 2*NL-1 barriers and 3 passes over W (5 with momentum) per iteration, against the
// reference's ~3*NL+4 device-wide syncs and 5 (9) passes.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "kp_common.cuh"

namespace hpnn {

// TG: tagged-pair hand-over between ranks (world > 1); false = plain buffers + barriers.
// grid / bid: CTAs of this rank and this CTA's index among them (the whole launch, except in
// the virtual-rank launch below where several ranks share one cooperative grid).
template <bool TG>
__device__ __forceinline__ void train_sample_body(const KP &p, const int grid, const int bid) {
  extern __shared__ double smem[];
  watch_init(p);
  if (p.fault == 1) return;  // injected fault: this rank never shows up
  double *xs = smem;
  double *ds = xs + p.max_ld;
  double *red = ds + p.max_rows;
  double *rowpart = red + kRedDoubles;

  const hpnn_dev_net &n = p.net;
  const int NL = n.n_layers;
  const int n_out = n.dims[NL];
  double *A = n.arena[n.rank];
  const bool snn = p.type == HPNN_TYPE_SNN;
  const int upd = (p.train == HPNN_TRAIN_BPM) ? 2 : 1;

  Barrier bar;
  bar.init(n, grid, bid);
  // Several ranks: every vector that crosses NVLink is handed over as tagged pairs (no fences,
  // no cross-rank barrier); grid barriers stay local to the GPU.  p.tagged == 0 keeps the
  // fence + flag-word barrier (one rank, or HPNN_TP_PROTOCOL=barrier).
  constexpr bool tg = TG;
  const unsigned long long launch = (unsigned long long)ld_relaxed_gpu(n.local_sync + 160) + 1ull;
  auto tag_of = [&](int iteration) { return (launch << 32) | (unsigned long long)(unsigned)iteration; };
  auto sync_phase = [&](bool remote) {
    if (tg) bar.sync_local();
    else bar.sync_all(n, remote);
  };
  int trace_n = 0;
  auto trace = [&](int tag) {
    if (p.trace && bid == 0 && threadIdx.x == 0 && trace_n < p.trace_cap) {
      p.trace[2 * trace_n] = tag;
      p.trace[2 * trace_n + 1] = clock64();
      ++trace_n;
    }
  };

  // act[i], i >= 1, is double buffered: the fused update+forward of layer 0 rewrites
  // act[1] while other CTAs may still need the old values to finish delta_0.
  int cur = 0;
  auto act_of = [&](int rank, int i, int buf) -> double * {  // tagged mode: a buffer holds pairs
    return n.arena[rank] + n.off_act[i] + ((i == 0) ? 0 : (size_t)buf * pad8(n.dims[i]) * (tg ? 2 : 1));
  };
  // input vector of layer l (act[l], buffer src_buf) into shared memory
  auto stage_act = [&](double *dst_smem, int i, int buf, unsigned long long tag) {
    if (tg && i > 0) stage_pairs<false>(dst_smem, act_of(n.rank, i, buf), n.dims[i], pad8(n.dims[i]), tag);
    else stage_vec(dst_smem, act_of(n.rank, i, buf), pad8(n.dims[i]));
  };
  auto forward_layer = [&](int l, int src_buf, int dst_buf, int u, unsigned long long tag) {
    stage_act(xs, l, src_buf, tag);
    __syncthreads();
    Dst dst;
    dst.n = n.world;
    dst.tag = tg ? tag : 0ull;
    dst.mc = (!tg && n.arena_mc) ? n.arena_mc + (act_of(n.rank, l + 1, dst_buf) - A) + n.row0[l] : nullptr;
    for (int q = 0; q < n.world; ++q) dst.p[q] = act_of(q, l + 1, dst_buf) + (size_t)n.row0[l] * (tg ? 2 : 1);
    const int kind = (snn && l == NL - 1) ? 1 : 0;
    const int segs = pick_row_segments(n.rows[l], n.ld[l], grid);
    if (u == 0) phase_rows<0>(n.W[l], nullptr, n.ld[l], n.rows[l], xs, ds, 0.0, 0.0, kind, dst, segs, rowpart, grid, bid);
    else if (u == 1) phase_rows<1>(n.W[l], nullptr, n.ld[l], n.rows[l], xs, ds, p.lr, p.alpha, kind, dst, segs, rowpart, grid, bid);
    else phase_rows<2>(n.W[l], n.dW[l], n.ld[l], n.rows[l], xs, ds, p.lr, p.alpha, kind, dst, segs, rowpart, grid, bid);
  };

  // ---- prologue: zero the momentum (per sample, src/ann.c:2386), first forward
  if (p.mode == 0 && upd == 2) {
    for (int l = 0; l < NL; ++l) {
      double2 *m2 = reinterpret_cast<double2 *>(n.dW[l]);
      const size_t tot = (size_t)n.rows[l] * n.ld[l] / 2;
      for (size_t i = (size_t)bid * blockDim.x + threadIdx.x; i < tot; i += (size_t)grid * blockDim.x)
        stg_cg2(m2 + i, make_double2(0.0, 0.0));
    }
  }
  for (int l = 0; l < NL; ++l) {
    forward_layer(l, cur, cur, 0, tag_of(0));
    sync_phase(bid < n.rows[l]);
  }
  const double *target = A + n.off_target;
  const int last = NL - 1;
  // output statistics from the gathered output vector (tagged mode: gathered into xs first)
  auto evaluate = [&](unsigned long long tag, int row0, int rows, double *dsp, double *o_out) -> OutStats {
    if (tg) {
      __syncthreads();
      stage_act(xs, NL, cur, tag);
      __syncthreads();
      return output_stats<true>(p.type, p.stable, xs, target, n_out, row0, rows, dsp, o_out, red);
    }
    return output_stats(p.type, p.stable, act_of(n.rank, NL, cur), target, n_out, row0, rows, dsp, o_out, red);
  };
  OutStats st = evaluate(tag_of(0), n.row0[last], n.rows[last], ds,
                         (p.mode == 1 && bid == 0) ? A + n.off_out : nullptr);
  double E = st.err;
  const double init_err = E;
  int iter = 0, ok = (st.argmax == st.target), first_ok = 0;
  double dE = 0.0;

  if (p.mode == 0) {
    for (;;) {
      ++iter;
      trace(1);
      const unsigned long long abort_seen = watch_peek();
      const unsigned long long tag = tag_of(iter);
      // ---- top-down: deltas from the old weights fused with the update
      for (int l = last; l >= 1; --l) {
        if (l != last) {
          stage_delta(n, l, layer_plan(n, l + 1, grid).rsplit, act_of(n.rank, l + 1, cur), ds, tg, tag);
          __syncthreads();
        }
        const int cols = n.dims[l];
        const ColPlan plan = layer_plan(n, l, grid);
        const int cb = plan.cb;
        const long long off_prev = n.off_part[l - 1];
        const int stride_prev = n.part_stride[l - 1];
        auto make_sink = [&](int c0, int rp) {
          return [=, &n](int i, double v) {
            const int col = c0 + i;
            if (col < cols) {
              const int own = owner_of_row(col, cols, n.world);
              const size_t slot = (size_t)(n.rank * kMaxRsplit + rp) * stride_prev + (col - row0_of(own, cols, n.world));
              if (tg) st_pair<true>(n.arena[own] + off_prev, (long long)slot, v, tag);
              else stg_cg(n.arena[own] + off_prev + slot, v);
            }
          };
        };
        const double *h = act_of(n.rank, l, cur);
        if (tg) {  // the input vector of layer l from the previous sweep: present, but stored as pairs
          stage_act(xs, l, cur, tag - 1);
          __syncthreads();
          h = xs;
        }
        if (upd == 1) phase_cols<1>(n.W[l], nullptr, n.ld[l], n.rows[l], cb, plan.rsplit, ds, h, tg, p.lr, p.alpha, red, make_sink, grid, bid);
        else phase_cols<2>(n.W[l], n.dW[l], n.ld[l], n.rows[l], cb, plan.rsplit, ds, h, tg, p.lr, p.alpha, red, make_sink, grid, bid);
        trace(100 + l);
        sync_phase(bid < (n.ld[l] / cb) * plan.rsplit);
        trace(200 + l);
      }
      // ---- bottom-up: layer 0 update fused with its forward, then the rest
      stage_delta(n, 0, layer_plan(n, 1, grid).rsplit, act_of(n.rank, 1, cur), ds, tg, tag);
      forward_layer(0, 0, cur ^ 1, upd, tag);
      trace(300);
      sync_phase(bid < n.rows[0]);
      trace(400);
      for (int l = 1; l < NL; ++l) {
        forward_layer(l, cur ^ 1, cur ^ 1, 0, tag);
        trace(300 + l);
        sync_phase(bid < n.rows[l]);
        trace(400 + l);
      }
      cur ^= 1;
      // ---- evaluate (every CTA, identical arithmetic => identical decision)
      watch_publish(abort_seen);
      st = evaluate(tag, n.row0[last], n.rows[last], ds, nullptr);
      trace(2);
      if (s_watch.exit_now) break;  // the job was abandoned (a wait ran out of time)
      dE = E - st.err;
      E = st.err;
      ok = (st.argmax == st.target);
      if (iter == 1) first_ok = ok;
      if (iter > p.max_iter) break;
      ok = ok && (iter > p.min_iter);
      if (!(dE > p.thr || !ok)) break;
    }
    // publish the final (normalised) output vector
    if (bid == 0) evaluate(tag_of(iter), 0, 0, nullptr, A + n.off_out);
  }
  bar.finish();
  if (bid == 0 && threadIdx.x == 0) n.local_sync[160] = (unsigned)launch;
  if (bid == 0 && threadIdx.x == 0 && p.result) {
    hpnn_train_result r;
    r.init_err = init_err;
    r.final_de = dE;
    r.last_err = E;
    r.n_iter = iter;
    r.first_ok = first_ok;
    r.success = ok;
    r.status = watch_status();
    *p.result = r;
  }
}

template <bool TG>
__global__ void __launch_bounds__(kThreads, 1) k_train_sample(const __grid_constant__ KP p) {
  train_sample_body<TG>(p, (int)gridDim.x, (int)blockIdx.x);
}

// Virtual ranks: ONE cooperative launch hosts `world` ranks of a row-split job on one GPU, each
// with its own weights slice, arena and sync words and gridDim.x / world CTAs.  It exists so that
// the multi-rank protocols (peer stores, cross-rank barrier, tagged hand-over) can be verified
// where only one GPU is available -- the successor of the reference's DEBUG build that fakes three
// GPUs with three contexts on device 0 (include/libhpnn/common.h:511-572).  Ranks that wait on each
// other must be co-resident, which separate launches on one GPU do not guarantee; a single
// cooperative grid does.
template <bool TG>
__global__ void __launch_bounds__(kThreads, 1) k_train_sample_vranks(const KP *pv, int ctas_per_rank) {
  __shared__ KP sp;
  const int vr = (int)blockIdx.x / ctas_per_rank, bid = (int)blockIdx.x % ctas_per_rank;
  {
    const unsigned long long *src = reinterpret_cast<const unsigned long long *>(pv + vr);
    unsigned long long *dst = reinterpret_cast<unsigned long long *>(&sp);
    static_assert(sizeof(KP) % 8 == 0, "KP is copied in 8-byte words");
    for (int i = threadIdx.x; i < (int)(sizeof(KP) / 8); i += kThreads) dst[i] = src[i];
  }
  __syncthreads();
  train_sample_body<TG>(sp, ctas_per_rank, bid);
}

// defined in train_resident.cu: 0 = launched, < 0 = error, 1 = not eligible (use the streaming kernel)
int launch_resident_if_eligible(const KP &p, int grid, void *stream);
// defined in train_lazy.cu, same convention
int launch_lazy_if_eligible(const KP &p, int grid, size_t smem, void *stream);
// defined in train_lazy1.cu (one GPU, row panels), same convention
int launch_lazy1_if_eligible(const KP &p, int grid, void *stream);
int prepare_lazy1(const KP &p_in, int grid, KP &p_out, size_t &smem_out);
int launch_lazy1_vranks(const KP *pv_dev, int world, int ctas_per_rank, bool tagged, size_t smem, void *stream);
extern long long g_lazy_min_bytes;

// ------------------------------- host side ---------------------------------
static int g_grid_override = 0;
static long long g_spin_timeout_ms = -1;  // -1: HPNN_SPIN_TIMEOUT_MS or 10 s
static int g_fault_kind = 0, g_fault_rank = -1;
static int g_jitter_ns = 0;
static long long *g_trace = nullptr;
static int g_trace_cap = 0;
static thread_local char g_err[512] = "";
long long g_launches = 0;

void set_error(const char *what, cudaError_t e) {
  snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
}
void set_error_msg(const char *what) { snprintf(g_err, sizeof(g_err), "%s", what); }

}  // namespace hpnn

using namespace hpnn;

extern "C" {

const char *hpnn_last_error(void) { return g_err; }

long long hpnn_dev_launch_count(int reset) {
  long long v = g_launches;
  if (reset) g_launches = 0;
  return v;
}

long long hpnn_abi_sizeof(int which) {
  switch (which) {
    case 0: return (long long)sizeof(hpnn_train_result);
    case 1: return (long long)sizeof(hpnn_train_opts);
    case 2: return (long long)sizeof(hpnn_conf);
    case 3: return (long long)sizeof(hpnn_dev_net);
    default: return -1;
  }
}

void hpnn_dev_set_persistent_grid(int ctas) { g_grid_override = ctas; }

void hpnn_dev_set_spin_timeout_ms(long long ms) { g_spin_timeout_ms = ms; }

void hpnn_dev_set_jitter_ns(int ns) { g_jitter_ns = ns > 0 ? ns : 0; }

void hpnn_dev_set_fault(int kind, int rank) {
  g_fault_kind = kind;
  g_fault_rank = rank;
}

void hpnn_dev_set_phase_trace(long long *dev_buf, int capacity_pairs) {
  g_trace = dev_buf;
  g_trace_cap = dev_buf ? capacity_pairs : 0;
}

int hpnn_dev_persistent_grid(void) {
  int dev = 0, sms = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (g_grid_override > 0 && g_grid_override <= sms) return g_grid_override;
  return sms;
}

long long hpnn_dev_arena_layout(hpnn_dev_net *net) {
  if (!net || net->n_layers < 2 || net->n_layers > HPNN_MAX_LAYERS || net->world < 1 ||
      net->world > HPNN_MAX_WORLD || net->rank < 0 || net->rank >= net->world)
    return -1;
  const int NL = net->n_layers, W = net->world;
  auto pad = [](long long v) { return (v + 7) & ~7LL; };
  long long off = 0;
  for (int l = 0; l < NL; ++l) {
    const int N = net->dims[l + 1];
    net->ld[l] = (int)pad(net->dims[l]);
    net->row0[l] = (int)((long long)net->rank * N / W);
    net->rows[l] = (int)((long long)(net->rank + 1) * N / W) - net->row0[l];
    net->part_stride[l] = (int)pad((N + W - 1) / W);
  }
  for (int i = 0; i <= NL; ++i) {
    net->off_act[i] = off;
    off += pad(net->dims[i]) * (i == 0 ? 1 : 4);  // two buffers; each may hold {value, tag} pairs
  }
  for (int l = 0; l < NL; ++l) {
    net->off_part[l] = off;
    off += 2LL * W * kMaxRsplit * net->part_stride[l];  // slot (rank, row part), {value, tag} pairs
  }
  net->off_target = off;
  off += pad(net->dims[NL]);
  net->off_out = off;
  off += pad(net->dims[NL]);
  net->off_flags = off;
  off += 2 * HPNN_MAX_WORLD;
  // per-CTA partial slots of the resident trainer (single rank, modest layers only)
  net->cta_slots = 0;
  for (int l = 0; l < NL; ++l) net->off_cta_part[l] = 0;
  if (W == 1) {
    long long cols = 0;
    for (int l = 1; l < NL; ++l) cols += net->ld[l];
    const int slots = 160;  // >= SM count of every Blackwell / Hopper part
    if (cols * slots * 16 <= (32LL << 20)) {
      net->cta_slots = slots;
      for (int l = 1; l < NL; ++l) {
        net->off_cta_part[l] = off;
        off += 2LL * slots * net->ld[l];  // {value, tag} pairs
      }
    }
  }
  // update-factor history of the deferred-update trainer: only where the weights of the layers
  // l >= 1 are far larger than L2, i.e. where a saved pass over W is worth the bookkeeping
  net->hist_depth = 0;
  for (int l = 0; l < NL; ++l) net->off_hist_u[l] = net->off_hist_v[l] = net->off_hist_dot[l] = 0;
  {
    long long wbytes = 0;
    for (int l = 1; l < NL; ++l) wbytes += 8LL * net->part_stride[l] * net->ld[l];
    if (wbytes >= (g_lazy_min_bytes >= 0 ? g_lazy_min_bytes : kLazyMinBytes)) {
      net->hist_depth = kLazyChunk;
      for (int l = 1; l < NL; ++l) {
        net->off_hist_u[l] = off;
        off += (long long)kLazyChunk * net->part_stride[l];
        net->off_hist_v[l] = off;
        off += (long long)kLazyChunk * net->ld[l];
        net->off_hist_dot[l] = off;
        off += 2LL * kLazyChunk * W * kLazyDotSlots;  // may hold {value, tag} pairs
      }
    }
  }
  net->off_panel = 0;
  net->panel_stride = 0;
  if (net->hist_depth > 0) {
    int wide = 8;
    for (int l = 1; l < NL; ++l) wide = net->ld[l] > wide ? net->ld[l] : wide;
    net->panel_stride = wide + kLazyChunk;
    net->off_panel = off;
    off += 2LL * kLazyDotSlots * net->panel_stride;
  }
  net->off_wlast = net->off_mlast = net->off_zpart = 0;
  if (net->hist_depth > 0 && net->dims[NL] <= kFoldMaxOut) {
    const long long wl = (long long)net->dims[NL] * net->ld[NL - 1];
    net->off_wlast = off;
    off += pad(wl);
    net->off_mlast = off;
    off += pad(wl);
    net->off_zpart = off;
    off += (long long)W * kLazyDotSlots * kFoldMaxOut;
  }
  net->arena_doubles = off;
  return off;
}

// Kernel parameters of one rank for a launch with `grid` CTAs per rank; smem_plain is the dynamic
// shared memory of the streaming kernel (k_train_sample).  0 = ok.
static int build_kp(const hpnn_dev_net *net, const hpnn_train_opts *o, int mode, hpnn_train_result *result_dev,
                    int grid, KP &p, size_t &smem_plain) {
  p.net = *net;
  hpnn_train_opts d;
  hpnn_train_opts_default(&d, o->type, o->train, 1);
  p.type = o->type;
  p.train = o->train;
  p.mode = mode;
  p.stable = o->stable_softmax;
  p.lr = o->lr > 0 ? o->lr : d.lr;
  p.alpha = o->alpha;
  p.thr = o->delta > 0 ? o->delta : d.delta;
  p.min_iter = o->min_iter >= 0 ? o->min_iter : d.min_iter;
  p.max_iter = o->max_iter >= 0 ? o->max_iter : d.max_iter;
  p.result = result_dev;
  {
    // Tagged hand-over for nets that stream < 48 MB per rank: their phases last a few microseconds and
    // a {value, tag} pair needs no fence at all (2x on SNN 784-1000-1000-10 at 4 GPUs).  Wide layers keep
    // plain buffers + the flag barrier: measured on 784-8192x4-10 the tagged variant pays ~10 k cycles
    // per phase for polling and unpacking 8192 pairs in every CTA (457 vs 405 us/iteration at 2 ranks,
    // profiles/r2_probe2b_wide.jsonl), and since the barrier got down to one fence.sys it costs ~9 k.
    const char *proto = getenv("HPNN_TP_PROTOCOL");
    long long local_doubles = 0;
    for (int l = 0; l < net->n_layers; ++l) local_doubles += (long long)net->rows[l] * net->ld[l];
    bool tagged = net->world > 1 && local_doubles * 8 < (48LL << 20);
    if (proto && !strcmp(proto, "barrier")) tagged = false;
    if (proto && !strcmp(proto, "tagged")) tagged = net->world > 1;
    p.tagged = tagged ? 1 : 0;
  }
  p.lazy_depth = 0;
  p.prefetch_bytes = 0;
  p.fold_out = 0;
  p.slicewise = 0;
  p.keep_l2 = 0;
  {
    static const long long env_ms = [] {
      const char *e = getenv("HPNN_SPIN_TIMEOUT_MS");
      return e ? atoll(e) : 10000LL;
    }();
    const long long ms = g_spin_timeout_ms >= 0 ? g_spin_timeout_ms : env_ms;
    p.spin_budget_ns = ms > 0 ? (unsigned long long)ms * 1000000ull : 0ull;
    p.fault = (g_fault_kind == 1 && g_fault_rank == net->rank) ? 1 : 0;
    p.jitter_ns = g_jitter_ns;
  }
  p.trace = g_trace;
  p.trace_cap = g_trace_cap;
  {
    static const int ns_env = [] {
      const char *e = getenv("HPNN_TRACE_NS");
      return (e && e[0] == '1') ? 1 : 0;
    }();
    p.trace_ns = ns_env;  // SM cycles follow the clock, which moves with the power draw of the phase before
  }
  int max_ld = 8, max_rows = 8;
  for (int l = 0; l < net->n_layers; ++l) {
    max_ld = net->ld[l] > max_ld ? net->ld[l] : max_ld;
    max_rows = net->rows[l] > max_rows ? net->rows[l] : max_rows;
    if (mode == 0 && o->train == HPNN_TRAIN_BPM && net->rows[l] > 0 && !net->dW[l]) {
      set_error_msg("hpnn_dev_train_sample: BPM needs momentum buffers");
      return -1;
    }
  }
  p.max_ld = max_ld;
  p.max_rows = (max_rows + 7) & ~7;
  int rowpart = 8;
  for (int l = 0; l < net->n_layers; ++l) {
    const int d = row_part_doubles(net->rows[l], grid);
    rowpart = d > rowpart ? d : rowpart;
  }
  smem_plain = sizeof(double) * ((size_t)p.max_ld + p.max_rows + kRedDoubles + rowpart);
  return 0;
}

int hpnn_dev_train_sample(const hpnn_dev_net *net, const hpnn_train_opts *o, int mode,
                          hpnn_train_result *result_dev, void *stream) {
  if (!net || !o) return -1;
  KP p;
  const int grid = hpnn_dev_persistent_grid();
  size_t smem = 0;
  if (build_kp(net, o, mode, result_dev, grid, p, smem) != 0) return -1;
  {
    const int rc = launch_lazy1_if_eligible(p, grid, stream);
    if (rc <= 0) {
      if (rc == 0) ++g_launches;
      return rc;
    }
  }
  if (smem <= kMaxDynSmem) {
    // BP on layers far larger than L2: deferred rank-1 updates save most of the write-back traffic
    const int rc = launch_lazy_if_eligible(p, grid, smem, stream);
    if (rc <= 0) {
      if (rc == 0) ++g_launches;
      return rc;
    }
  }
  {
    // small nets: keep every weight (and momentum) row in shared memory for the whole sample
    const int rc = launch_resident_if_eligible(p, grid, stream);
    if (rc <= 0) {
      if (rc == 0) ++g_launches;
      return rc;
    }
  }
  if (smem > kMaxDynSmem) {
    set_error_msg("hpnn_dev_train_sample: layer too wide for the persistent kernel (use the multi-launch path)");
    return -2;
  }
  const void *kfn = p.tagged ? (const void *)k_train_sample<true> : (const void *)k_train_sample<false>;
  cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) {
    set_error("cudaFuncSetAttribute(k_train_sample)", e);
    return -1;
  }
  void *args[] = {(void *)&p};
  e = cudaLaunchCooperativeKernel(kfn, dim3(grid), dim3(kThreads), args, smem,
                                  (cudaStream_t)stream);
  if (e != cudaSuccess) {
    set_error("cudaLaunchCooperativeKernel(k_train_sample)", e);
    return -1;
  }
  ++g_launches;
  return 0;
}

int hpnn_dev_train_sample_vranks(const hpnn_dev_net *nets, int world, const hpnn_train_opts *o, int mode,
                                 hpnn_train_result *const *results_dev, void *stream) {
  if (!nets || !o || world < 1 || world > HPNN_MAX_WORLD) return -1;
  int sms = hpnn_dev_persistent_grid();
  const int G = sms / world;
  if (G < 1) {
    set_error_msg("hpnn_dev_train_sample_vranks: more ranks than SMs");
    return -1;
  }
  KP kp[HPNN_MAX_WORLD];
  memset((void *)kp, 0, sizeof(kp));
  size_t smem = 0;
  int lazy1 = -1;
  for (int r = 0; r < world; ++r) {
    if (nets[r].world != world || nets[r].rank != r) {
      set_error_msg("hpnn_dev_train_sample_vranks: nets[r] must be rank r of `world`");
      return -1;
    }
    size_t sm_r = 0;
    KP base;
    memset((void *)&base, 0, sizeof(base));
    if (build_kp(nets + r, o, mode, results_dev ? results_dev[r] : nullptr, G, base, sm_r) != 0) return -1;
    if (r > 0) base.trace = nullptr;  // one trace buffer: rank 0 writes it
    KP pl;
    size_t sm_l = 0;
    const int el = prepare_lazy1(base, G, pl, sm_l) == 0 ? 1 : 0;
    if (lazy1 < 0) lazy1 = el;
    if (el != lazy1) {
      set_error_msg("hpnn_dev_train_sample_vranks: ranks disagree on the trainer variant");
      return -1;
    }
    kp[r] = el ? pl : base;
    const size_t s_r = el ? sm_l : sm_r;
    smem = s_r > smem ? s_r : smem;
    if (kp[r].tagged != kp[0].tagged) {
      set_error_msg("hpnn_dev_train_sample_vranks: ranks disagree on the protocol");
      return -1;
    }
  }
  if (smem + sizeof(KP) > kMaxDynSmem) {
    set_error_msg("hpnn_dev_train_sample_vranks: layer too wide for the virtual-rank launch");
    return -2;
  }
  // parameter block of every rank in device memory (one buffer per device, created on first use)
  static KP *pv_dev[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return -1;
  cudaError_t e = cudaSuccess;
  if (!pv_dev[dev]) e = cudaMalloc((void **)&pv_dev[dev], sizeof(KP) * HPNN_MAX_WORLD);
  if (e == cudaSuccess)
    e = cudaMemcpyAsync(pv_dev[dev], kp, sizeof(KP) * world, cudaMemcpyHostToDevice, (cudaStream_t)stream);
  if (e != cudaSuccess) {
    set_error("hpnn_dev_train_sample_vranks: parameter upload", e);
    return -1;
  }
  if (lazy1) {
    const int rc = launch_lazy1_vranks(pv_dev[dev], world, G, kp[0].tagged != 0, smem, stream);
    if (rc == 0) ++g_launches;
    return rc;
  }
  const void *kfn = kp[0].tagged ? (const void *)k_train_sample_vranks<true> : (const void *)k_train_sample_vranks<false>;
  e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) {
    set_error("cudaFuncSetAttribute(k_train_sample_vranks)", e);
    return -1;
  }
  const KP *pv = pv_dev[dev];
  int g = G;
  void *args[] = {(void *)&pv, (void *)&g};
  e = cudaLaunchCooperativeKernel(kfn, dim3(world * G), dim3(kThreads), args, smem, (cudaStream_t)stream);
  if (e != cudaSuccess) {
    set_error("cudaLaunchCooperativeKernel(k_train_sample_vranks)", e);
    return -1;
  }
  ++g_launches;
  return 0;
}

}  // extern "C"
