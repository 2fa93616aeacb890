// libgnnuds.so - host orchestration + C ABI (include/gnnuds.h) of the B200-native
// Emulator forward.  Mirrors surrogate/emulator.py::Emulator (build_network 166-341,
// predict 566-602, post_proc 643-678, constrain 727-748) of Zhiyu014/GNN-UDS without
// sharing code with it: the Keras graph becomes a fixed sequence of sm_100a kernels
// over [B][T][N][C] activations.
#include "../../include/gnnuds.h"
#include "kernels.cuh"
#include "tc_gemm.cuh"
#include "tc_selftest.cuh"
#include "fused_sp.cuh"
#include "fused_conv.cuh"
#include "fused_agg.cuh"
#include "fused_aggt.cuh"
#include "fused_tp.cuh"
#include "gru_tc.cuh"
#include "nn_cat.cuh"
#include "backward.cuh"

#include <algorithm>
#include <array>
#include <thread>
#include <cmath>
#include <map>
#include <mutex>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <set>
#include <string>
#include <type_traits>
#include <vector>

using namespace gu;

namespace {

thread_local std::string g_create_error;

struct LayerW { int t[4]; };   // tensor indices of one weighted layer

struct SpatialW {
  LayerW xe, ex, ne_n, ne_m, conv_x, conv_e;   // graph_base == 0
  LayerW joint;                                // graph_base != 0
  int cx, ce;
};

struct TensorDesc { int64_t numel; int d0, d1, d2; int last() const { return d2 ? d2 : (d1 ? d1 : d0); } };

// Activation tape of one fp32 forward chunk, kept for the reverse pass (csrc/backward.cuh, backward_host.inc): with the
// tape on, every layer whose output the backward needs (post-activation values for act', h = x@W of the GAT / GCN
// aggregations) writes into its own slice of a bump-allocated arena instead of the ping-pong workspace buffers.
struct SpTape { float *XE, *EX, *HBx, *HBe, *xo, *eo; };
struct Tape {
  bool on = false;
  float* base = nullptr; size_t cap = 0, used = 0;     // floats
  SpTape sp[2][16];
  float* tp[2][2][8];                                  // [block][0 node / 1 edge][layer] temporal layer outputs
  float* mx[2][2][8];                                  // GRU / LSTM: input projection x@K + b of every step
  float* hall[2][2][8];                                // GRU / LSTM: hidden state of every step
  float* call[2][2][8];                                // LSTM: cell state of every step
  float *resx = nullptr, *rese = nullptr;              // activated resnet outputs (the heads' input)
  float* flood[8];
  float* alloc(size_t n) {
    n = (n + 63) / 64 * 64;
    if (!on || used + n > cap) return nullptr;
    float* p = base + used; used += n; return p;
  }
};

}  // namespace

struct gnnuds_handle {
  gnnuds_config c{};
  std::string err;
  cudaStream_t stream = nullptr;
  int sm_count = 148;
  int64_t launches = 0;
  bool attr_done[8] = {false, false, false, false, false, false, false, false};   // per handle (= per device): opt-in to > 48 KB dynamic smem
  bool has_graph = false, has_weights = false, has_norm = false;
  int debug_stop = 0;
  int pp_variant = 0;           // 0 post_proc (numpy rules), 1 post_proc_tf (gnnuds_set_post_proc_variant); rollouts always use 1
  bool pump_all = false;        // every link is a pump (emulator.py:698 pump.min() > 0)
  int profiling = 0;            // 0 off, 1 whole-call event pair, 2 also one pair per GEMM launch
  float last_total_ms = 0.f, last_gemm_ms = 0.f;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> gemm_ev;
  std::vector<int> ev_family;           // kernel family of each event pair (GNNUDS_FAM_*)
  size_t gemm_ev_used = 0;
  float fam_ms[4] = {0.f, 0.f, 0.f, 0.f};
  int64_t fam_launches[4] = {0, 0, 0, 0};
  cudaEvent_t ev_a = nullptr, ev_b = nullptr;

  // layout
  std::vector<TensorDesc> tens;
  LayerW emb_x, emb_b, emb_e, emb_a;
  std::vector<SpatialW> sp[2];
  std::vector<LayerW> tpx[2], tpe[2];
  LayerW resx, rese, head_out, head_flood, head_e;
  std::vector<LayerW> flood_chain;
  // NN-cat (conv == GNNUDS_CONV_NONE): Dense(2E) spatial layers and zero-padded copies of the layers whose flattened
  // widths (N*C) are not multiples of 16 (csrc/nn_cat.cuh)
  std::vector<LayerW> nn_sp[2];
  bool nn() const { return c.conv == GNNUDS_CONV_NONE; }
  float* d_nnpad = nullptr;
  struct NnPad { long long w = 0, b = 0; int K = 0, Kp = 0, No = 0, Nop = 0; };
  NnPad nnp_x, nnp_b, nnp_e, nnp_a, nnp_out, nnp_flood, nnp_eout;

  // device parameters
  float* d_blob = nullptr;
  std::vector<int64_t> toff;
  uint8_t* d_wimg = nullptr;            // bf16 K-major weight images of every GEMM layer (tensor-core path)
  std::vector<long long> img_off;       // byte offset per tensor index, -1 = no image
  bool tma_on = true;                   // GNNUDS_NO_TMA=1 forces the cp.async producers (diagnostics)
  bool tc() const { return c.precision == GNNUDS_PREC_BF16_TC; }
  size_t es() const { return tc() ? 2 : 4; }   // bytes per stored activation element
  std::vector<float*> d_ne_matT[2];     // per spatial layer (block*S + l): N-side, M-side folded matrices
  int ldm_n = 0, ldm_m = 0, ti_n = 0, ti_m = 0;
  // bf16 K-major images of the folded NodeEdge matrices for the fused spatial kernel (side 0: N x M, 1: M x N)
  std::vector<uint8_t*> d_ne_img[2];
  uint32_t ne_img_bytes[2] = {0, 0}, ne_pstride[2] = {0, 0};
  int ne_kp[2] = {0, 0};
  // small graphs: sp_conv_kernel processes sp_g consecutive (b,t) samples as ONE block-diagonal super-sample
  // (sp_g*R rows fill the 128-row MMA tile: NodeEdge matrix and CSR are replicated on the diagonal)
  int sp_g = 1;
  int gru_tc = 1;                       // GNNUDS_GRU_TC: 0 = fp32 SIMT recurrence (gru_kernel) on the tensor-core path
  int sw128 = 1;                        // GNNUDS_SW128: 0 = 16-byte panel boxes for every TMA-fed operand tile
  int conv1d_tma = 1;                   // GNNUDS_CONV1D_TMA: 0 = cp.async gathers for time-shifted segments
  int fuse_heads = 1;                   // GNNUDS_FUSE_HEADS: 0 = separate cumsum / head kernels (debugging)
  int fused_mode = 1;                   // GNNUDS_FUSED: 0 layer-by-layer kernels, 1 fused spatial kernels (default)

  // graph
  int *d_rp = nullptr, *d_ci = nullptr, *d_erp = nullptr, *d_eci = nullptr;
  int* d_bins[2] = {nullptr, nullptr};  // sp_conv gather schedule per stream side: bin_ptr[SP_BINS+1] | bin_rows[R] | row_len[R]
  // CSR padded to 4 entries per block (zero-weight pads) for the 16-byte index / weight loads of sp_conv_kernel
  int* d_prp[2] = {nullptr, nullptr}; int* d_pci[2] = {nullptr, nullptr}; float* d_pval[2] = {nullptr, nullptr};
  int pnnz[2] = {0, 0};
  int sp_max_len[2] = {0, 0};           // longest true CSR row per side (selects sp_conv's register softmax)
  // sp_agg_kernel (csrc/fused_agg.cuh): unpadded block-diagonal CSR with per-entry alpha offsets, GCN values
  int* d_arp[2] = {nullptr, nullptr}; uint32_t* d_acj[2] = {nullptr, nullptr}; float* d_aval[2] = {nullptr, nullptr};
  int annz[2] = {0, 0};
  std::vector<uint8_t*> d_cv_img[2];    // GAT conv weight images with the score columns (NW = E + 16 rows per panel)
  std::vector<uint32_t> cv_img_bytes[2];
  // use_adj (GAT): per-entry controlled-link index / static keep flags of the node filter, the chunk's settings
  uint8_t* d_zeros = nullptr;           // 64 KB of zeros: source of sp_agg_kernel's bulk re-zeroing of its alpha region
  int* d_adj_act = nullptr; unsigned char* d_adj_keep = nullptr; bool use_adj = false; const float* cur_a = nullptr;
  int sp_aggt = 1;                      // GNNUDS_SP_AGGT: 0 = sp_agg_kernel (rows on M, two tiles) for 128 < R <= 144 instead of sp_aggt_kernel
  std::vector<uint8_t*> d_wa[2];        // GAT: score operand images per conv layer and side (sp_aggt_kernel, fused_aggt.cuh)
  int tp_fused = 1;                     // GNNUDS_TP_FUSED: 0 = one tc_rows_kernel launch per Conv1D layer instead of tp_conv_kernel
  int sp_agg = 1;                       // GNNUDS_SP_AGG: 0 = sp_conv_kernel (SIMT gather) instead of sp_agg_kernel
  float *d_val = nullptr, *d_eval = nullptr;
  int nnz = 0, ennz = 0, n_adj = 0;
  int *d_edges = nullptr, *d_act_of_edge = nullptr, *d_act_out = nullptr, *d_act_in = nullptr;
  int *d_inc_ptr = nullptr, *d_inc_edge = nullptr, *d_inc_sign = nullptr;
  float *d_hmin = nullptr, *d_hmax = nullptr, *d_outfall = nullptr, *d_ehmax = nullptr;
  // pumped / offset networks (null on the others)
  float *d_pump = nullptr, *d_offset = nullptr, *d_pump_in = nullptr, *d_pump_out = nullptr, *d_area = nullptr;
  int *d_edge_from = nullptr, *d_ps = nullptr;
  int n_ps = 0;                         // pumped-storage nodes (emulator.py:591)
  bool has_pump = false, has_offset = false;
  std::vector<int> h_edges;

  // norms (max,min) per item
  float *d_nx = nullptr, *d_nb = nullptr, *d_ny = nullptr, *d_ne = nullptr;
  int cx_n = 0, cb_n = 0, cy_n = 0, ce_n = 0;

  // workspace
  int64_t ws_chunk = 0;
  float* d_ws = nullptr;
  size_t ws_floats = 0;
  float *XA, *XB, *EA, *EB, *HB, *XE, *EX, *NX, *NE, *TX0, *TX1, *TE0, *TE1, *BEMB, *AEMB, *RES, *RESE,
      *OUTN, *OUTE, *SS, *SN, *HBJ = nullptr, *J0 = nullptr, *J1 = nullptr;
  float *in_x, *in_b, *in_a, *in_ex, *out_y, *out_ey;      // chunk staging on device
  float *in2_x, *in2_b, *in2_a, *in2_ex, *out2_y, *out2_ey; // second staging set: copies of chunk i+1 / i-1 overlap compute of chunk i
  cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
  cudaEvent_t ev_in[2] = {nullptr, nullptr}, ev_comp[2] = {nullptr, nullptr}, ev_out[2] = {nullptr, nullptr};
  // pageable caller buffers (plain NumPy arrays): a ring of two pinned input / output sets; worker threads copy
  // between caller memory and the ring while the GPU computes (gnnuds_predict)
  float* pin_in[2] = {nullptr, nullptr}; float* pin_out[2] = {nullptr, nullptr};
  size_t pin_in_floats = 0, pin_out_floats = 0;
  // rollout state
  float *ro_x = nullptr, *ro_ex = nullptr, *ro_b = nullptr, *ro_a = nullptr;
  int64_t ro_B = 0;
  // grow-only device buffers of the autoregressive / MPC entry points (no cudaMalloc on the hot path once warm):
  // 0..3 rollout horizon b / a / y / ey, 4..10 objective (weights x3, gamma, y, q0, obj), 11..17 evaluate
  // (x0, ex0, b, dec, partial sums, prev_q, obj)
  float* gb[48] = {nullptr}; size_t gb_cap[48] = {0};   // 18..21 replay (raw states, perfs, edge states, settings)

  // CUDA graphs of launch-bound call bodies (small populations: the reverse pass issues ~3 k kernels per 12-chunk gradient).
  // A graph is keyed by the call's scalars, its caller pointers and buf_gen, which every (re)allocation of a handle buffer
  // bumps (the graphs bake device addresses in).  GNNUDS_GRAPHS=0 disables them.
  uint64_t buf_gen = 0;
  int use_graphs = 1;
  // Small batches: the node-side and edge-side kernels of a layer are independent and neither fills the GPU, so they run
  // concurrently - node side on `stream`, edge side on `s_side`, fork / join by events (par2).  Decided per chunk
  // (side_prepare); large batches keep the single-stream order.  GNNUDS_SIDE_PAR=0 disables it.
  cudaStream_t s_side = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  int side_par = 1;
  // fp32 GEMMs of at most this many rows go to the one-round 16 x 32 tile kernels (rows_gemm_f32_skinny); GNNUDS_SKINNY_ROWS=0 disables
  long long skinny_rows = 8192;
  int pdl = 1;      // small-tile GEMMs are launched with programmatic stream serialization (weight prefetch under the previous kernel's tail); GNNUDS_PDL=0 disables
  bool side_on = false;
  int default_ws_mb = 8192;             // workspace budget when cfg.workspace_mb == 0 (GNNUDS_WORKSPACE_MB overrides)
  int joint_per_sample = 0;             // GNNUDS_JOINT_PER_SAMPLE=1: one (b,t) sample at a time in the large joint conv (diagnostic: no gain at C5)
  float* HB2 = nullptr;                 // the edge side's projection / conv scratch while both sides are in flight
  std::vector<std::pair<std::string, cudaGraphExec_t>> graphs;
  int64_t n_capture = 0, n_replay = 0;
  std::set<std::string> graph_seen;                   // keys run once (eagerly): captured when they come back
  std::map<std::string, int64_t> graph_launches;      // kernels per graph (gnnuds_launch_count stays a kernel count)
  Tape tape;                            // reverse pass (gnnuds_predict_vjp / gnnuds_grad_settings)
  float* bw_arena = nullptr; size_t bw_cap = 0;

  // live streams for debug_fetch
  const float* dbg_ptr[2] = {nullptr, nullptr};
  int64_t dbg_rows[2] = {0, 0};
  int dbg_width[2] = {0, 0};

  int ny() const { return 3 + (c.if_flood ? 1 : 0) + 1; }
  int ldn() const { return c.n_out + (c.if_flood ? 1 : 0); }
};

namespace {

int fail(gnnuds_handle* h, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (h) h->err = buf; else g_create_error = buf;
  return code;
}

#define CU(call)                                                                          \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess)                                                                \
      return fail(h, GNNUDS_E_CUDA, "%s -> %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)

// ---- bf16 operand images for the tensor-core kernels (layout: csrc/umma.cuh) -------------
#if GU_F16
#define GU_TMAP_16 CU_TENSOR_MAP_DATA_TYPE_FLOAT16
#else
#define GU_TMAP_16 CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
#endif
inline uint16_t f2bf(float f) {          // float -> the build's 16-bit storage type (round to nearest even)
#if GU_F16
  return __half_as_ushort(__float2half_rn(f));
#else
  uint32_t u;
  std::memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fc0;
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
#endif
}
inline float bf2f(uint16_t b) {          // and back
#if GU_F16
  return __half2float(__ushort_as_half(b));
#else
  const uint32_t u = (uint32_t)b << 16;
  float f; std::memcpy(&f, &u, 4);
  return f;
#endif
}
// K-major panel image of mat[rows][cols] (element (r,k) = mat[r*ldr + k*ldk]); rows/cols beyond the
// matrix are zero.  Byte offset of (r,k): (k/8)*pstride + r*16 + (k%8)*2.
std::vector<uint16_t> pack_kmajor(const float* mat, int rows, int cols, long ldr, long ldk, int Rp, int Kp, int pstride) {
  std::vector<uint16_t> img((size_t)(Kp / 8) * pstride / 2, 0);
  for (int r = 0; r < rows; ++r)
    for (int k = 0; k < cols; ++k)
      img[((size_t)(k / 8) * pstride + (size_t)r * 16) / 2 + (k % 8)] = f2bf(mat[r * ldr + k * ldk]);
  (void)Rp;
  return img;
}

int conv_ntens(const gnnuds_config& c) { return c.conv == GNNUDS_CONV_GAT ? 4 : 2; }

// ---- creation-order walk of build_network (emulator.py:195-336) ----------------------
struct LayoutBuilder {
  gnnuds_handle* h;
  int add(int64_t n, int d0 = 0, int d1 = 0, int d2 = 0) {
    h->tens.push_back({n, d0, d1, d2});
    return (int)h->tens.size() - 1;
  }
  LayerW dense(int cin, int cout) {
    LayerW l{};
    l.t[0] = add((int64_t)cin * cout, cin, cout);
    l.t[1] = add(cout, cout);
    return l;
  }
  LayerW conv(int cin) {
    const int E = h->c.embed;
    LayerW l{};
    if (h->c.conv == GNNUDS_CONV_GAT) {
      l.t[0] = add((int64_t)cin * E, cin, 1, E);
      l.t[1] = add(E, E, 1, 1);
      l.t[2] = add(E, E, 1, 1);
      l.t[3] = add(E, E);
    } else {
      l.t[0] = add((int64_t)cin * E, cin, E);
      l.t[1] = add(E, E);
    }
    return l;
  }
  LayerW node_edge(int r, int cdim) {
    LayerW l{};
    l.t[0] = add((int64_t)r * cdim, r, cdim);
    l.t[1] = add((int64_t)r * cdim, r, cdim);
    return l;
  }
  void temporal(std::vector<LayerW>& v, int cin) {
    const auto& c = h->c;
    const int H = c.hidden;
    for (int i = 0; i < c.n_tp_layer; ++i) {
      const int ci = i == 0 ? cin : H;
      LayerW l{};
      if (c.recurrent == GNNUDS_REC_CONV1D) {
        l.t[0] = add((int64_t)c.kernel_size * ci * H, c.kernel_size, ci, H);
        l.t[1] = add(H, H);
      } else if (c.recurrent == GNNUDS_REC_GRU) {
        l.t[0] = add((int64_t)ci * 3 * H, ci, 3 * H);
        l.t[1] = add((int64_t)H * 3 * H, H, 3 * H);
        l.t[2] = add(2 * 3 * H, 2, 3 * H);
      } else {
        l.t[0] = add((int64_t)ci * 4 * H, ci, 4 * H);
        l.t[1] = add((int64_t)H * 4 * H, H, 4 * H);
        l.t[2] = add(4 * H, 4 * H);
      }
      v.push_back(l);
    }
  }
  void spatial(std::vector<SpatialW>& v, int cx, int ce) {
    const auto& c = h->c;
    const int E = c.embed;
    for (int l = 0; l < c.n_sp_layer; ++l) {
      SpatialW s{};
      s.cx = cx; s.ce = ce;
      if (c.graph_base) {
        s.joint = conv(cx);
      } else {
        s.xe = dense(ce, E / 2);                   // emulator.py:225
        s.ex = dense(cx, E / 2);                   // :226
        s.ne_n = node_edge(c.n_node, c.n_edge);    // :227
        s.ne_m = node_edge(c.n_edge, c.n_node);    // :228
        s.conv_x = conv(cx + E / 2);               // :229
        s.conv_e = conv(ce + E / 2);               // :230
      }
      v.push_back(s);
      cx = ce = E;
    }
  }
  // NN-cat creation order (emulator.py:195-336 with conv falsy)
  void build_nn() {
    const auto& c = h->c;
    const int E = c.embed, H = c.hidden, N = c.n_node, M = c.n_edge;
    h->emb_x = dense(N * c.n_in, E);               // :197-198
    h->emb_b = dense(N * c.b_in, E / 2);           // :202-203
    h->emb_e = dense(M * c.e_in, E);               // :205-206
    if (c.act) h->emb_a = dense(M, E / 2);         // :211-212
    for (int l = 0; l < c.n_sp_layer; ++l) h->nn_sp[0].push_back(dense(2 * E, 2 * E));   // :232
    temporal(h->tpx[0], E);
    temporal(h->tpe[0], E);
    int ci = 2 * H + E / 2 + (c.act ? E / 2 : 0);
    for (int l = 0; l < c.n_sp_layer; ++l) { h->nn_sp[1].push_back(dense(ci, 2 * E)); ci = 2 * E; }   // :285
    temporal(h->tpx[1], E);
    temporal(h->tpe[1], E);
    h->resx = dense(H, E);
    h->rese = dense(H, E);
    h->head_out = dense(E, c.n_out * N);           // :322-324
    if (c.if_flood) {
      int cf = E;
      for (int i = 0; i < c.if_flood; ++i) { h->flood_chain.push_back(dense(cf, E / 2)); cf = E / 2; }
      h->head_flood = dense(cf, N);                // :327,330
    }
    h->head_e = dense(E, c.e_out * M);             // :335-336
  }
  void build() {
    const auto& c = h->c;
    const int E = c.embed, H = c.hidden;
    if (h->nn()) { build_nn(); return; }
    h->emb_x = dense(c.n_in, E);                   // :198
    h->emb_b = dense(c.b_in, E / 2);               // :203
    h->emb_e = dense(c.e_in, E);                   // :206
    if (c.act) h->emb_a = dense(1, E / 2);         // :212
    spatial(h->sp[0], E, E);
    temporal(h->tpx[0], E);                        // :247
    temporal(h->tpe[0], E);                        // :254
    spatial(h->sp[1], H + E / 2, H + (c.act ? E / 2 : 0));
    temporal(h->tpx[1], E);                        // :302
    temporal(h->tpe[1], E);                        // :308
    h->resx = dense(H, E);                         // :313 'dense_resx'
    h->rese = dense(H, E);                         // :317
    h->head_out = dense(E, c.n_out);               // :324
    if (c.if_flood) {
      int ci = E;
      for (int i = 0; i < c.if_flood; ++i) { h->flood_chain.push_back(dense(ci, E / 2)); ci = E / 2; }   // :329
      h->head_flood = dense(ci, 1);                // :330
    }
    h->head_e = dense(E, c.e_out);                 // :336
  }
};

const float* W(const gnnuds_handle* h, int tensor) { return h->d_blob + h->toff[tensor]; }

int grid_for(const gnnuds_handle* h, long long work_items, int per_block, int max_waves = 8) {
  long long g = (work_items + per_block - 1) / per_block;
  long long cap = (long long)h->sm_count * max_waves;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

// ---- launch helpers -------------------------------------------------------------------

// per-launch CUDA event pairs (profiling level 2), tagged with a kernel family
int prof_begin(gnnuds_handle* h, int family, cudaEvent_t* e1) {
  *e1 = nullptr;
  if (h->profiling < 2) return 0;
  if (h->gemm_ev_used == h->gemm_ev.size()) {
    cudaEvent_t a, b;
    CU(cudaEventCreate(&a)); CU(cudaEventCreate(&b));
    h->gemm_ev.push_back({a, b});
    h->ev_family.push_back(0);
  }
  h->ev_family[h->gemm_ev_used] = family;
  *e1 = h->gemm_ev[h->gemm_ev_used].second;
  CU(cudaEventRecord(h->gemm_ev[h->gemm_ev_used].first, h->stream));
  ++h->gemm_ev_used;
  return 0;
}
int prof_end(gnnuds_handle* h, cudaEvent_t e1) {
  if (e1) CU(cudaEventRecord(e1, h->stream));
  return 0;
}

// Node side (fx) on h->stream and edge side (fe) on h->s_side, concurrently when side_on; joined before returning.
// Inside fe, h->stream IS the side stream and h->HB the side's own scratch, so every launch helper works unchanged.
template <class FX, class FE>
int par2(gnnuds_handle* h, FX fx, FE fe) {
  if (!h->side_on) { const int rc = fx(); return rc ? rc : fe(); }
  cudaStream_t main = h->stream;
  float* hb = h->HB;
  CU(cudaEventRecord(h->ev_fork, main));
  CU(cudaStreamWaitEvent(h->s_side, h->ev_fork, 0));
  const int rc1 = fx();
  h->stream = h->s_side; h->HB = h->HB2;
  const int rc2 = rc1 ? 0 : fe();
  h->stream = main; h->HB = hb;
  CU(cudaEventRecord(h->ev_join, h->s_side));
  CU(cudaStreamWaitEvent(main, h->ev_join, 0));
  return rc1 ? rc1 : rc2;
}

// Run `body` (stream-ordered work on h->stream only: kernel launches, async memsets / D2D copies on handle or caller
// device buffers; no allocation, no synchronisation) as a CUDA graph cached under `key`.  First use captures and
// instantiates; a capture that fails (something in the body is not capturable) switches graphs off for the handle and
// the body runs eagerly.  The caller has made every buffer the body touches exist BEFORE this call.
template <class F>
int run_graphed(gnnuds_handle* h, std::string key, F body) {
  if (!h->use_graphs || h->profiling >= 2 || h->debug_stop) return body();
  char gen[48];
  snprintf(gen, sizeof gen, "|g%llu|v%d", (unsigned long long)h->buf_gen, h->pp_variant);
  key += gen;
  cudaGraphExec_t exec = nullptr;
  for (auto& g : h->graphs) if (g.first == key) { exec = g.second; break; }
  if (!exec) {
    // a shape is captured on its SECOND use: capture + instantiation cost more than one eager run, and callers such as the
    // mode-B event replay see every (length, chunk count) once per pass
    if (h->graph_seen.insert(key).second) {
      if (h->graph_seen.size() > 4096) h->graph_seen.clear();
      return body();
    }
    if (cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeRelaxed) != cudaSuccess) { (void)cudaGetLastError(); h->use_graphs = 0; return body(); }
    const int64_t l0 = h->launches;
    const int rc = body();
    cudaGraph_t g = nullptr;
    const cudaError_t e = cudaStreamEndCapture(h->stream, &g);
    if (rc || e != cudaSuccess || !g) {
      (void)cudaGetLastError();
      if (g) cudaGraphDestroy(g);
      if (rc) return rc;
      h->use_graphs = 0; h->launches = l0;
      return body();
    }
    if (cudaGraphInstantiate(&exec, g, 0) != cudaSuccess) { (void)cudaGetLastError(); cudaGraphDestroy(g); h->use_graphs = 0; h->launches = l0; return body(); }
    cudaGraphDestroy(g);
    if (h->graphs.size() >= 256) { cudaGraphExecDestroy(h->graphs.front().second); h->graph_launches.erase(h->graphs.front().first); h->graphs.erase(h->graphs.begin()); }
    h->graphs.push_back({key, exec});
    h->graph_launches[key] = h->launches - l0;
    // a caller that keeps producing new shapes pays for captures it never replays: give up on graphs for this handle
    if (++h->n_capture > 128 && h->n_replay < 2 * h->n_capture) h->use_graphs = 0;
  } else {
    h->launches += h->graph_launches[key];
    ++h->n_replay;
  }
  CU(cudaGraphLaunch(exec, h->stream));
  return 0;
}

struct SegH { const float* p; int ld; int width; int shift; int pitch; int off; };

// Launch of a tensor-core kernel whose set-up (barrier init, TMEM allocation, constant staging) ends in tc::pdl_sync():
// small grids (at most half the SMs: single events, small MPC populations) are launched as programmatic dependents, so that
// set-up runs under the tail of the previous kernel of the stream; large grids are plain stream launches.
template <typename K, typename P>
cudaError_t launch_pdl(gnnuds_handle* h, K kernel, dim3 grid, unsigned threads, size_t smem, const P& p) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = smem; cfg.stream = h->stream;
  cudaLaunchAttribute at{};
  at.id = cudaLaunchAttributeProgrammaticStreamSerialization; at.val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = &at;
  cfg.numAttrs = (h->pdl && (long long)grid.x * grid.y * grid.z * 2 <= h->sm_count) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, p);
}

int launch_gemm(gnnuds_handle* h, const std::vector<SegH>& segs, const float* Wp, const float* bias, float* out,
                int ldo, int Nout, int act, long long Bc, int Nn, int Td, int Ts, int pitch_d, int off_d) {
  if (segs.size() > 4) return fail(h, GNNUDS_E_INVALID, "rows_gemm: more than 4 segments (kernel_size > 4?)");
  GemmP p{};
  p.nseg = (int)segs.size();
  for (int i = 0; i < p.nseg; ++i) {
    if (segs[i].width % 16) return fail(h, GNNUDS_E_INVALID, "rows_gemm: segment width %d not a multiple of 16", segs[i].width);
    p.seg[i] = Seg{segs[i].p, segs[i].ld, segs[i].width, segs[i].shift, segs[i].pitch, segs[i].off};
  }
  p.W = Wp; p.bias = bias; p.out = out; p.ldo = ldo; p.Nout = Nout; p.act = act;
  p.rows = Bc * Td * Nn; p.Nn = Nn; p.Td = Td; p.Ts = Ts; p.pitch_d = pitch_d; p.off_d = off_d;
  if (p.rows == 0) return 0;
  cudaEvent_t e1 = nullptr;
  { int prc = prof_begin(h, GNNUDS_FAM_GEMM, &e1); if (prc) return prc; }
  const unsigned gx = (unsigned)((p.rows + 127) / 128);
  int Ktot = 0;
  for (int i = 0; i < p.nseg; ++i) Ktot += p.seg[i].width;
  const size_t skinny_smem = ((size_t)16 * (Ktot + 4) + (size_t)32 * Ktot) * sizeof(float);
  if (p.rows <= h->skinny_rows && Nout % 32 == 0 && skinny_smem <= 200 * 1024) {
    if (!h->attr_done[7]) {
      CU(cudaFuncSetAttribute(rows_gemm_f32_skinny, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      h->attr_done[7] = true;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)((p.rows + 15) / 16), Nout / 32); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = skinny_smem; cfg.stream = h->stream;
    cudaLaunchAttribute at{};
    at.id = cudaLaunchAttributeProgrammaticStreamSerialization; at.val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = &at; cfg.numAttrs = h->pdl ? 1 : 0;
    CU(cudaLaunchKernelEx(&cfg, rows_gemm_f32_skinny, p, Ktot));
  } else if (Nout % 128 == 0) {
    rows_gemm_f32<128><<<dim3(gx, Nout / 128), 256, 0, h->stream>>>(p);
  } else if (Nout % 64 == 0) {
    rows_gemm_f32<64><<<dim3(gx, Nout / 64), 256, 0, h->stream>>>(p);
  } else if (Nout % 32 == 0) {
    rows_gemm_f32_narrow<32><<<dim3(gx, Nout / 32), 256, 0, h->stream>>>(p);
  } else if (Nout % 16 == 0) {
    rows_gemm_f32_narrow<16><<<dim3(gx, Nout / 16), 256, 0, h->stream>>>(p);
  } else {
    return fail(h, GNNUDS_E_INVALID, "rows_gemm: Nout %d not a multiple of 16", Nout);
  }
  ++h->launches;
  { int prc = prof_end(h, e1); if (prc) return prc; }
  CU(cudaGetLastError());
  return 0;
}

template <typename ST = float>
int launch_embed(gnnuds_handle* h, const float* in, const float* nrm, int cn, int Nn_norm, const int* gather, int n_act,
                 const LayerW& w, void* out, float* res, long long Bc, int T, int Nn, int K, int Eo, int act, int mode) {
  EmbedP p{};
  p.in = in; p.nmax = nrm; p.nmin = nrm ? nrm + (size_t)Nn_norm * cn : nullptr; p.cn = cn;
  p.gather = gather; p.n_act = n_act;
  p.W = W(h, w.t[0]); p.bias = W(h, w.t[1]); p.out = out; p.res = res;
  p.rows = Bc * T * Nn; p.T = T; p.Nn = Nn; p.K = K; p.Eo = Eo; p.act = act; p.mode = mode;
  if (p.rows == 0) return 0;
  if (std::is_same<ST, bf16>::value && mode == 0 && K <= 8 && (Eo == 128 || Eo == 64) && p.rows < (1ll << 31)) {
    const int g16 = grid_for(h, p.rows, 16, 8), g8 = grid_for(h, p.rows, 32, 8);
#define GU_EMBED(KT) do { if (Eo == 128) embed_bf16_kernel<16, KT><<<g16, 256, 0, h->stream>>>(p); \
                          else embed_bf16_kernel<8, KT><<<g8, 256, 0, h->stream>>>(p); } while (0)
    switch (K) { case 1: GU_EMBED(1); break; case 2: GU_EMBED(2); break; case 4: GU_EMBED(4); break; case 5: GU_EMBED(5); break;
                 default: GU_EMBED(0); }
#undef GU_EMBED
    ++h->launches;
    CU(cudaGetLastError());
    return 0;
  }
  const int rpb = 256 / (Eo / 4);
  embed_kernel<ST><<<grid_for(h, p.rows, rpb, 16), 256, (size_t)(K * Eo + Eo) * sizeof(float), h->stream>>>(p);
  ++h->launches;
  CU(cudaGetLastError());
  return 0;
}

template <int TI, typename T>
int launch_ne_t(gnnuds_handle* h, const NodeEdgeP& p, size_t smem) {
  CU(cudaFuncSetAttribute(nodeedge_kernel<TI, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int g = grid_for(h, (p.S + 1) / 2, 1, 1);
  nodeedge_kernel<TI, T><<<g, 512, smem, h->stream>>>(p);
  ++h->launches;
  CU(cudaGetLastError());
  return 0;
}

template <typename T = float>
int launch_nodeedge(gnnuds_handle* h, const float* matT, int ldm, int TI, const void* in, void* out, long long S,
                    int R_in, int R_out, int F) {
  NodeEdgeP p{matT, in, out, S, R_in, R_out, F, ldm};
  const size_t smem = ((size_t)R_in * ldm + 2 * (size_t)R_in * F) * sizeof(float);
  if (smem > 227 * 1024) return fail(h, GNNUDS_E_INVALID, "NodeEdge matrix (%d x %d) does not fit shared memory", R_out, R_in);
  switch (TI) {
    case 4: return launch_ne_t<4, T>(h, p, smem);
    case 8: return launch_ne_t<8, T>(h, p, smem);
    case 12: return launch_ne_t<12, T>(h, p, smem);
    case 16: return launch_ne_t<16, T>(h, p, smem);
    default: return fail(h, GNNUDS_E_INVALID, "NodeEdge: rows per thread %d unsupported (graph too large for the NodeEdge path)", TI);
  }
}

template <typename T = float>
int launch_agg(gnnuds_handle* h, const void* hin, void* out, const int* rp, const int* ci, const float* val,
               const LayerW& w, long long S, int R, int act, bool dyn = false) {
  const auto& c = h->c;
  const int E = c.embed;
  const bool gat = c.conv == GNNUDS_CONV_GAT;
  const float* bias = W(h, w.t[gat ? 3 : 1]);
  const size_t smem = ((size_t)R * E + 2 * (size_t)R) * sizeof(float);
  if (smem <= 200 * 1024) {
    AggP p{hin, out, rp, ci, val, gat ? W(h, w.t[1]) : nullptr, gat ? W(h, w.t[2]) : nullptr, bias, S, R, E, act, gat ? 1 : 0,
           nullptr, nullptr, nullptr, 0};
    if (dyn && h->use_adj && h->cur_a) { p.act_entry = h->d_adj_act; p.keep = h->d_adj_keep; p.a_dyn = h->cur_a; p.n_act = c.n_act; }
    CU(cudaFuncSetAttribute(agg_small_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int threads = S <= 2 * (long long)h->sm_count ? 1024 : 256;      // few samples: one warp per graph row at once
    agg_small_kernel<T><<<grid_for(h, S, 1, 16), threads, smem, h->stream>>>(p);
    ++h->launches;
  } else {
    if (dyn && h->use_adj) return fail(h, GNNUDS_E_INVALID, "use_adj: graph of %d rows does not fit the per-sample aggregation kernel", R);
    if (gat) {
      row_scores_kernel<T><<<grid_for(h, S * R, 8, 16), 256, 0, h->stream>>>(static_cast<const T*>(hin), W(h, w.t[1]), W(h, w.t[2]), h->SS, h->SN, S * R, E);
      ++h->launches;
    }
    AggLargeP p{hin, out, rp, ci, val, h->SS, h->SN, bias, S, R, E, act, gat ? 1 : 0};
    cudaEvent_t e1 = nullptr;
    { int prc = prof_begin(h, GNNUDS_FAM_GATHER, &e1); if (prc) return prc; }
    const int agg16 = getenv("GNNUDS_AGG16") ? atoi(getenv("GNNUDS_AGG16")) : 1;
    if (std::is_same<T, bf16>::value && E == 128 && agg16 == 2) {       // 16-byte gathers, one 128-byte row half per pass (diagnostic)
      if (gat) agg_large16_kernel<8, true, 4><<<grid_for(h, (S * R + 3) / 4, 8, 32), 256, 0, h->stream>>>(p);
      else agg_large16_kernel<8, false, 4><<<grid_for(h, (S * R + 3) / 4, 8, 32), 256, 0, h->stream>>>(p);
    } else if (std::is_same<T, bf16>::value && E == 128 && agg16 == 1) {  // 16-byte gathers, whole rows, half a warp per row
      if (gat) agg_large16_kernel<16, true, 4><<<grid_for(h, (S * R + 1) / 2, 8, 32), 256, 0, h->stream>>>(p);
      else if (getenv("GNNUDS_AGG_MINB") && atoi(getenv("GNNUDS_AGG_MINB")) == 4) agg_large16_kernel<16, false, 4><<<grid_for(h, (S * R + 1) / 2, 8, 32), 256, 0, h->stream>>>(p);
      else agg_large16_kernel<16, false, 5><<<grid_for(h, (S * R + 1) / 2, 8, 40), 256, 0, h->stream>>>(p);   // 5 CTAs per SM: more rows in flight
    } else
      agg_large_kernel<T><<<grid_for(h, S * R, 8, 32), 256, 0, h->stream>>>(p);
    ++h->launches;
    { int prc = prof_end(h, e1); if (prc) return prc; }
  }
  CU(cudaGetLastError());
  return 0;
}

template <typename T = float>
int launch_head(gnnuds_handle* h, const void* in, int ld, int K, const LayerW& w, float* out, int ldo, int col0,
                int Nout, int act, long long rows) {
  if (Nout > 4) return fail(h, GNNUDS_E_INVALID, "head wider than 4 columns (%d)", Nout);
  HeadP p{in, ld, K, W(h, w.t[0]), W(h, w.t[1]), out, ldo, col0, Nout, act, rows};
  if (rows == 0) return 0;
  head_kernel<T><<<grid_for(h, rows, 8, 16), 256, (size_t)K * Nout * sizeof(float), h->stream>>>(p);
  ++h->launches;
  CU(cudaGetLastError());
  return 0;
}

int act_code(const gnnuds_config& c) { return c.activation; }

// ---- temporal stack (emulator.py:154-163) ------------------------------------------------
// in: (Bc,Ts,Nn,cin) at `src`; result (Bc,Td,Nn,H) in *dst (one of t0/t1).
int run_temporal(gnnuds_handle* h, const std::vector<LayerW>& L, const SegH& src, int cin, long long Bc, int Nn,
                 int Ts, int Td, float* t0, float* t1, float** dst, float** tape_slots = nullptr) {
  const auto& c = h->c;
  const int H = c.hidden, P = c.n_tp_layer;
  const float* cur = src.p;
  int cw = cin, cld = src.ld, cpitch = src.pitch, coff = src.off;
  float* bufs[2] = {t0, t1};
  for (int i = 0; i < P; ++i) {
    const bool last = (i == P - 1);
    const int Tdst = last ? Td : Ts;
    float* o = bufs[i & 1];
    if (h->tape.on && tape_slots) {
      tape_slots[i] = o = h->tape.alloc((size_t)Bc * Tdst * Nn * H);
      if (!o) return fail(h, GNNUDS_E_NOMEM, "activation tape exhausted (temporal layer)");
    }
    int rc;
    if (c.recurrent == GNNUDS_REC_CONV1D) {
      std::vector<SegH> segs;
      const int k = c.kernel_size, d = 1 << i;
      for (int j = 0; j < k; ++j) segs.push_back({cur, cld, cw, (k - 1 - j) * d, cpitch, coff});
      rc = launch_gemm(h, segs, W(h, L[i].t[0]), W(h, L[i].t[1]), o, H, H, act_code(c), Bc, Nn, Tdst, Ts, Nn, 0);
      if (rc) return rc;
    } else {
      const bool gru = c.recurrent == GNNUDS_REC_GRU;
      const int G = (gru ? 3 : 4) * H;
      float *mxb = h->HB, *hall = nullptr, *call = nullptr;
      if (h->tape.on && tape_slots) {            // reverse pass (BPTT): the projection and every step's state stay on the tape
        const long idx = (long)(tape_slots - &h->tape.tp[0][0][0]) / 8;
        const size_t rws = (size_t)Bc * Ts * Nn;
        h->tape.mx[idx / 2][idx % 2][i] = mxb = h->tape.alloc(rws * G);
        h->tape.hall[idx / 2][idx % 2][i] = hall = h->tape.alloc(rws * H);
        if (!gru) h->tape.call[idx / 2][idx % 2][i] = call = h->tape.alloc(rws * H);
        if (!mxb || !hall || (!gru && !call)) return fail(h, GNNUDS_E_NOMEM, "activation tape exhausted (recurrent layer)");
      }
      // input projection for every time step
      rc = launch_gemm(h, {{cur, cld, cw, 0, cpitch, coff}}, W(h, L[i].t[0]), W(h, L[i].t[2]) /* b[0] or b */, mxb, G, G,
                       ACT_LINEAR, Bc, Nn, Ts, Ts, Nn, 0);
      if (rc) return rc;
      const long long nseq = Bc * Nn;
      const int g = (int)((nseq + 31) / 32);
      if (gru) {
        GruP p{mxb, W(h, L[i].t[1]), W(h, L[i].t[2]) + G, o, nseq, Nn, Ts, Tdst, H, hall};
        const size_t smem = ((size_t)H * G + 32 * H + 32 * G) * sizeof(float);
        CU(cudaFuncSetAttribute(gru_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        cudaEvent_t e1 = nullptr;
        { int prc = prof_begin(h, GNNUDS_FAM_RECURRENT, &e1); if (prc) return prc; }
        gru_kernel<float><<<g, 256, smem, h->stream>>>(p);
        { int prc = prof_end(h, e1); if (prc) return prc; }
      } else {
        LstmP p{mxb, W(h, L[i].t[1]), o, nseq, Nn, Ts, Tdst, H, hall, call};
        const size_t smem = ((size_t)H * G + 64 * H + 32 * G) * sizeof(float);
        CU(cudaFuncSetAttribute(lstm_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        cudaEvent_t e1 = nullptr;
        { int prc = prof_begin(h, GNNUDS_FAM_RECURRENT, &e1); if (prc) return prc; }
        lstm_kernel<float><<<g, 256, smem, h->stream>>>(p);
        { int prc = prof_end(h, e1); if (prc) return prc; }
      }
      ++h->launches;
      CU(cudaGetLastError());
    }
    cur = o; cw = H; cld = H; cpitch = Nn; coff = 0;
    // note: for non-final layers Tdst == Ts so the next layer reads a full-length sequence
  }
  *dst = const_cast<float*>(cur);
  return 0;
}

struct Stream2 { std::vector<SegH> x, e; };   // current node / edge stream as concatenated segments
int side_prepare(gnnuds_handle* h, int64_t Bc);
template <bool TC> int run_spatial_joint(gnnuds_handle* h, int blk, Stream2& st, long long Bc, int T);

// ---- spatial block (emulator.py:219-235 / 272-288), graph_base == 0 -----------------------
int run_spatial(gnnuds_handle* h, int blk, Stream2& st, long long Bc, int T) {
  const auto& c = h->c;
  const int E = c.embed, E2 = E / 2, N = c.n_node, M = c.n_edge;
  const long long S = Bc * T;
  const int act = act_code(c);
  const int nt = conv_ntens(c);
  (void)nt;
  for (int l = 0; l < c.n_sp_layer; ++l) {
    const SpatialW& w = h->sp[blk][l];
    int rc;
    float *XE = h->XE, *EX = h->EX, *HBx = h->HB, *HBe = h->HB;
    float* xo = (l & 1) ? h->XB : h->XA;
    float* eo = (l & 1) ? h->EB : h->EA;
    if (h->tape.on) {            // reverse pass: these outputs stay alive on the tape
      SpTape& tp = h->tape.sp[blk][l];
      tp.XE = XE = h->tape.alloc((size_t)S * M * E2); tp.EX = EX = h->tape.alloc((size_t)S * N * E2);
      tp.HBx = HBx = h->tape.alloc((size_t)S * N * E); tp.HBe = HBe = h->tape.alloc((size_t)S * M * E);
      tp.xo = xo = h->tape.alloc((size_t)S * N * E); tp.eo = eo = h->tape.alloc((size_t)S * M * E);
      if (!XE || !EX || !HBx || !HBe || !xo || !eo) return fail(h, GNNUDS_E_NOMEM, "activation tape exhausted (spatial layer)");
    }
    // x_e = Dense(E/2)(e) ; e_x = Dense(E/2)(x)   (the node side consumes x_e, the edge side e_x)
    if ((rc = par2(h, [&] { return launch_gemm(h, st.x, W(h, w.ex.t[0]), W(h, w.ex.t[1]), EX, E2, E2, act, Bc, N, T, T, N, 0); },
                   [&] { return launch_gemm(h, st.e, W(h, w.xe.t[0]), W(h, w.xe.t[1]), XE, E2, E2, act, Bc, M, T, T, M, 0); }))) return rc;
    const int li = blk * c.n_sp_layer + l;
    const bool taped = h->tape.on;
    // per side: NodeEdge of the other side's Dense, conv on [self | NodeEdge], aggregation
    auto node_side = [&]() -> int {
      int r;
      if ((r = launch_nodeedge(h, h->d_ne_matT[0][li], h->ldm_n, h->ti_n, XE, h->NX, S, M, N, E2))) return r;
      std::vector<SegH> sx = st.x; sx.push_back({h->NX, E2, E2, 0, N, 0});
      float* hb = taped ? HBx : h->HB;
      if ((r = launch_gemm(h, sx, W(h, w.conv_x.t[0]), nullptr, hb, E, E, ACT_LINEAR, Bc, N, T, T, N, 0))) return r;
      return launch_agg(h, hb, xo, h->d_rp, h->d_ci, h->d_val, w.conv_x, S, N, act, blk == 1);
    };
    auto edge_side = [&]() -> int {
      int r;
      if ((r = launch_nodeedge(h, h->d_ne_matT[1][li], h->ldm_m, h->ti_m, EX, h->NE, S, N, M, E2))) return r;
      std::vector<SegH> se = st.e; se.push_back({h->NE, E2, E2, 0, M, 0});
      float* hb = taped ? HBe : h->HB;                   // inside par2's edge branch h->HB is the side's own scratch
      if ((r = launch_gemm(h, se, W(h, w.conv_e.t[0]), nullptr, hb, E, E, ACT_LINEAR, Bc, M, T, T, M, 0))) return r;
      return launch_agg(h, hb, eo, h->d_erp, h->d_eci, h->d_eval, w.conv_e, S, M, act);
    };
    if ((rc = par2(h, node_side, edge_side))) return rc;
    st.x = {{xo, E, E, 0, N, 0}};
    st.e = {{eo, E, E, 0, M, 0}};
    if (h->debug_stop == (blk ? 40 : 20) + l) return 1;
  }
  return 0;
}

void set_dbg(gnnuds_handle* h, const Stream2& st, long long rows_x, long long rows_e) {
  const bool joint = st.x[0].pitch != h->c.n_node;   // joint node+edge buffer: no compact view to fetch
  h->dbg_ptr[0] = joint ? nullptr : st.x[0].p; h->dbg_width[0] = st.x[0].width; h->dbg_rows[0] = rows_x;
  h->dbg_ptr[1] = joint ? nullptr : st.e[0].p; h->dbg_width[1] = st.e[0].width; h->dbg_rows[1] = rows_e;
}

int run_postproc(gnnuds_handle* h, const float* x, const float* b, const float* a, long long Bc, float* y, float* ey,
                 bool normalised_io) {
  const auto& c = h->c;
  const int N = c.n_node, M = c.n_edge, To = c.seq_out;
  const int ldn = h->ldn();
  // ---- P1..P4 post-processing ----
  PostP p{};
  p.out_n = h->OUTN; p.ldn = ldn; p.out_e = h->OUTE; p.lde = c.e_out;
  p.b_raw = b; p.a = c.act ? a : nullptr; p.n_act = c.n_act;
  p.act_of_edge = h->d_act_of_edge; p.act_out_of_node = h->d_act_out; p.act_in_of_node = h->d_act_in;
  p.inc_ptr = h->d_inc_ptr; p.inc_edge = h->d_inc_edge; p.inc_sign = h->d_inc_sign;
  p.ymax = h->d_ny; p.ymin = h->d_ny + (size_t)N * h->cy_n; p.cy = h->cy_n;
  p.emax = h->d_ne; p.emin = h->d_ne + (size_t)M * h->ce_n; p.ce = h->ce_n;
  p.bmax = h->d_nb; p.bmin = h->d_nb + (size_t)N * h->cb_n; p.cb = h->cb_n;
  p.hmin = h->d_hmin; p.hmax = h->d_hmax; p.is_outfall = h->d_outfall; p.ehmax = h->d_ehmax;
  p.y = y; p.ey = ey; p.S = Bc * To; p.N = N; p.M = M; p.b_in = c.b_in; p.n_out = c.n_out; p.e_out = c.e_out;
  p.ny = normalised_io ? 3 + (c.if_flood ? 1 : 0) : h->ny();
  p.if_flood = c.if_flood ? 1 : 0; p.edge_fusion = c.edge_fusion; p.act = c.act; p.tide = c.tide;
  p.normalised = normalised_io ? 1 : 0; p.epsilon = c.epsilon; p.tf_rules = h->pp_variant;
  if (normalised_io && c.tide) return fail(h, GNNUDS_E_INVALID, "ROLL_NORM with tide not supported");
  const bool big = (size_t)(M + ((h->has_pump || h->has_offset) ? N : 0)) * sizeof(float) > 40 * 1024;      // large graphs: edge flows staged in HBM scratch (SS covers Bc*T*max(N,M))
  p.efl_g = big ? h->SS : nullptr;
  const bool need_h = h->has_pump || h->has_offset;
  if (need_h) {
    p.edge_from = h->d_edge_from; p.pump = (h->has_pump && (!h->pp_variant || h->pump_all)) ? h->d_pump : nullptr; p.offset = h->has_offset ? h->d_offset : nullptr;
    p.pump_in = h->has_pump ? h->d_pump_in : nullptr; p.pump_out = h->has_pump ? h->d_pump_out : nullptr;
    p.hn_g = big ? h->SN : nullptr;
  }
  const size_t psm = big ? 0 : (size_t)(M + (need_h ? N : 0)) * sizeof(float);
  postproc_kernel<<<grid_for(h, p.S, 1, 16), big ? 512 : 128, psm, h->stream>>>(p);
  ++h->launches;
  CU(cudaGetLastError());
  if (h->n_ps && !normalised_io) {
    // P3: pumped-storage depth integrated over the horizon (emulator.py:590-598), then the h-dependent part of constrain
    PumpStoreP q{y, h->ny(), x, c.n_in, b, c.b_in, h->d_ps, h->n_ps, h->d_area, h->d_hmin, h->d_hmax, h->d_outfall,
                 Bc, c.seq_in, To, N, c.if_flood ? 1 : 0, c.epsilon};
    const long long work = Bc * h->n_ps;
    pump_storage_kernel<<<(unsigned)((work + 127) / 128), 128, 0, h->stream>>>(q);
    ++h->launches;
    CU(cudaGetLastError());
  }
  return 0;
}

// One forward over Bc scenarios whose inputs are already on the device.
// normalised_io: inputs already normalised, outputs left normalised (ROLL_NORM rule).
int forward_chunk(gnnuds_handle* h, const float* x, const float* b, const float* a, const float* ex, long long Bc,
                  float* y, float* ey, bool normalised_io) {
  const auto& c = h->c;
  const int E = c.embed, E2 = E / 2, H = c.hidden, N = c.n_node, M = c.n_edge, Ti = c.seq_in, To = c.seq_out;
  const int act = act_code(c);
  int rc;
  if ((rc = side_prepare(h, Bc))) return rc;
  // ---- K1 embedding ----
  if ((rc = par2(h,
      [&]() -> int {
        int r;
        if ((r = launch_embed(h, x, normalised_io ? nullptr : h->d_nx, h->cx_n, N, nullptr, 0, h->emb_x, h->XB, h->RES, Bc, Ti, N,
                              c.n_in, E, act, 0))) return r;
        return launch_embed(h, b, normalised_io ? nullptr : h->d_nb, h->cb_n, N, nullptr, 0, h->emb_b, h->BEMB, nullptr, Bc, To, N,
                            c.b_in, E2, act, 0);
      },
      [&]() -> int {
        int r;
        if ((r = launch_embed(h, ex, normalised_io ? nullptr : h->d_ne, h->ce_n, M, nullptr, 0, h->emb_e, h->EB, h->RESE, Bc, Ti, M,
                              c.e_in, E, act, 0))) return r;
        return c.act ? launch_embed(h, a, nullptr, 0, 0, h->d_act_of_edge, c.n_act, h->emb_a, h->AEMB, nullptr, Bc, To, M, 1, E2, act, 2) : 0;
      }))) return rc;
  Stream2 st;
  st.x = {{h->XB, E, E, 0, N, 0}};
  st.e = {{h->EB, E, E, 0, M, 0}};
  set_dbg(h, st, Bc * Ti * N, Bc * Ti * M);
  if (h->debug_stop == 1) return 0;
  // the first spatial layer writes XA/EA (l even), so the embedding lives in XB/EB
  // ---- spatial block 1 ----
  rc = c.graph_base ? run_spatial_joint<false>(h, 0, st, Bc, Ti) : run_spatial(h, 0, st, Bc, Ti);
  set_dbg(h, st, Bc * Ti * N, Bc * Ti * M);
  if (rc == 1) return 0;
  if (rc) return rc;
  if (h->debug_stop == 2) return 0;
  // ---- temporal block 1 (keeps the last To steps, emulator.py:249,256) ----
  float *xt = nullptr, *et = nullptr;
  if ((rc = par2(h, [&] { return run_temporal(h, h->tpx[0], st.x[0], E, Bc, N, Ti, To, h->TX0, h->TX1, &xt, h->tape.tp[0][0]); },
                 [&] { return run_temporal(h, h->tpe[0], st.e[0], E, Bc, M, Ti, To, h->TE0, h->TE1, &et, h->tape.tp[0][1]); }))) return rc;
  // ---- boundary / control concat (emulator.py:260-262) ----
  st.x = {{xt, H, H, 0, N, 0}, {h->BEMB, E2, E2, 0, N, 0}};
  st.e = {{et, H, H, 0, M, 0}};
  if (c.act) st.e.push_back({h->AEMB, E2, E2, 0, M, 0});
  set_dbg(h, st, Bc * To * N, Bc * To * M);
  if (h->debug_stop == 3) return 0;
  // ---- spatial block 2 ----
  rc = c.graph_base ? run_spatial_joint<false>(h, 1, st, Bc, To) : run_spatial(h, 1, st, Bc, To);
  set_dbg(h, st, Bc * To * N, Bc * To * M);
  if (rc == 1) return 0;
  if (rc) return rc;
  if (h->debug_stop == 4) return 0;
  // ---- temporal block 2 ----
  if ((rc = par2(h, [&] { return run_temporal(h, h->tpx[1], st.x[0], E, Bc, N, To, To, h->TX0, h->TX1, &xt, h->tape.tp[1][0]); },
                 [&] { return run_temporal(h, h->tpe[1], st.e[0], E, Bc, M, To, To, h->TE0, h->TE1, &et, h->tape.tp[1][1]); }))) return rc;
  st.x = {{xt, H, H, 0, N, 0}};
  st.e = {{et, H, H, 0, M, 0}};
  set_dbg(h, st, Bc * To * N, Bc * To * M);
  if (h->debug_stop == 5) return 0;
  // ---- K8 resnet + heads (emulator.py:312-338) ----
  const int ract = c.resnet ? ACT_LINEAR : act;
  float *XAo = h->XA, *EAo = h->EA;
  if (h->tape.on) {
    h->tape.resx = XAo = h->tape.alloc((size_t)Bc * To * N * E); h->tape.rese = EAo = h->tape.alloc((size_t)Bc * To * M * E);
    if (!XAo || !EAo) return fail(h, GNNUDS_E_NOMEM, "activation tape exhausted (resnet)");
  }
  const long long RXo = Bc * To * N, REo = Bc * To * M;
  const int ldn = h->ldn();
  // node side: Dense, resnet cumsum, hard_sigmoid head, flood chain + sigmoid head; edge side: Dense, cumsum, tanh head
  auto node_tail = [&]() -> int {
    int r;
    if ((r = launch_gemm(h, st.x, W(h, h->resx.t[0]), W(h, h->resx.t[1]), XAo, E, E, ract, Bc, N, To, To, N, 0))) return r;
    if (c.resnet) {
      cumsum_res_act_kernel<float><<<grid_for(h, Bc * N * (E / 4), 256, 16), 256, 0, h->stream>>>(XAo, h->RES, Bc, To, N, E, act);
      ++h->launches;
      CU(cudaGetLastError());
    }
    if (h->debug_stop == 6) return 0;
    if ((r = launch_head(h, XAo, E, E, h->head_out, h->OUTN, ldn, 0, c.n_out, ACT_HARD_SIGMOID, RXo))) return r;
    if (c.if_flood) {
      std::vector<SegH> cur = {{XAo, E, E, 0, N, 0}};
      float* bufs[2] = {h->EX, h->NX};
      for (int i = 0; i < c.if_flood; ++i) {
        float* fo = bufs[i & 1];
        if (h->tape.on) {
          if (i >= 8 || !(h->tape.flood[i] = fo = h->tape.alloc((size_t)RXo * E2))) return fail(h, GNNUDS_E_NOMEM, "activation tape exhausted (flood chain)");
        }
        if ((r = launch_gemm(h, cur, W(h, h->flood_chain[i].t[0]), W(h, h->flood_chain[i].t[1]), fo, E2, E2, act, Bc, N,
                             To, To, N, 0))) return r;
        cur = {{fo, E2, E2, 0, N, 0}};
      }
      if ((r = launch_head(h, cur[0].p, E2, E2, h->head_flood, h->OUTN, ldn, c.n_out, 1, ACT_SIGMOID, RXo))) return r;
    }
    return 0;
  };
  auto edge_tail = [&]() -> int {
    int r;
    if ((r = launch_gemm(h, st.e, W(h, h->rese.t[0]), W(h, h->rese.t[1]), EAo, E, E, ract, Bc, M, To, To, M, 0))) return r;
    if (c.resnet) {
      cumsum_res_act_kernel<float><<<grid_for(h, Bc * M * (E / 4), 256, 16), 256, 0, h->stream>>>(EAo, h->RESE, Bc, To, M, E, act);
      ++h->launches;
      CU(cudaGetLastError());
    }
    if (h->debug_stop == 6) return 0;
    return launch_head(h, EAo, E, E, h->head_e, h->OUTE, c.e_out, 0, c.e_out, ACT_TANH, REo);
  };
  if ((rc = par2(h, node_tail, edge_tail))) return rc;
  st.x = {{XAo, E, E, 0, N, 0}};
  st.e = {{EAo, E, E, 0, M, 0}};
  set_dbg(h, st, Bc * To * N, Bc * To * M);
  if (h->debug_stop == 6) return 0;
  return run_postproc(h, x, b, a, Bc, y, ey, normalised_io);
}

// ---- NN-cat forward (conv falsy): flattened rows through Dense / temporal stacks, csrc/nn_cat.cuh ----
int forward_chunk_nn(gnnuds_handle* h, const float* x, const float* b, const float* a, const float* ex, long long Bc,
                     float* y, float* ey, bool normalised_io) {
  const auto& c = h->c;
  const int E = c.embed, E2 = E / 2, H = c.hidden, N = c.n_node, M = c.n_edge, Ti = c.seq_in, To = c.seq_out;
  const int act = act_code(c);
  const float* P = h->d_nnpad;
  int rc;
  auto prep = [&](const float* in, const float* nrm, int cn, int Nn, int K, int Kp, int mode, long long rows) -> int {
    NnPrepP p{in, nrm, nrm ? nrm + (size_t)Nn * cn : nullptr, cn, h->d_act_of_edge, c.n_act, h->HB, rows, Nn, K, Kp, mode};
    nn_prep_kernel<<<grid_for(h, rows * Kp, 256, 16), 256, 0, h->stream>>>(p);
    ++h->launches;
    CU(cudaGetLastError());
    return 0;
  };
  auto unflat = [&](const float* in, int ldi, float* out, int ldo, int col0, long long rows, int Nn, int nc) -> int {
    NnUnflatP p{in, ldi, out, ldo, col0, rows, Nn, nc};
    nn_unflatten_kernel<<<grid_for(h, rows * Nn * nc, 256, 16), 256, 0, h->stream>>>(p);
    ++h->launches;
    CU(cudaGetLastError());
    return 0;
  };
  // ---- embeddings (emulator.py:195-213): x and e keep the pre-activation of the last input step for the resnet ----
  const float* nx = normalised_io ? nullptr : h->d_nx;
  const float* nb = normalised_io ? nullptr : h->d_nb;
  const float* ne = normalised_io ? nullptr : h->d_ne;
  if ((rc = prep(x, nx, h->cx_n, N, c.n_in, h->nnp_x.Kp, 0, Bc * Ti))) return rc;
  {
    const std::vector<SegH> s = {{h->HB, h->nnp_x.Kp, h->nnp_x.Kp, 0, 1, 0}};
    if ((rc = launch_gemm(h, s, P + h->nnp_x.w, W(h, h->emb_x.t[1]), h->XB, E, E, act, Bc, 1, Ti, Ti, 1, 0))) return rc;
    if ((rc = launch_gemm(h, s, P + h->nnp_x.w, W(h, h->emb_x.t[1]), h->RES, E, E, ACT_LINEAR, Bc, 1, 1, Ti, 1, 0))) return rc;
  }
  if ((rc = prep(b, nb, h->cb_n, N, c.b_in, h->nnp_b.Kp, 0, Bc * To))) return rc;
  if ((rc = launch_gemm(h, {{h->HB, h->nnp_b.Kp, h->nnp_b.Kp, 0, 1, 0}}, P + h->nnp_b.w, W(h, h->emb_b.t[1]), h->BEMB, E2, E2, act,
                        Bc, 1, To, To, 1, 0))) return rc;
  if ((rc = prep(ex, ne, h->ce_n, M, c.e_in, h->nnp_e.Kp, 0, Bc * Ti))) return rc;
  {
    const std::vector<SegH> s = {{h->HB, h->nnp_e.Kp, h->nnp_e.Kp, 0, 1, 0}};
    if ((rc = launch_gemm(h, s, P + h->nnp_e.w, W(h, h->emb_e.t[1]), h->EB, E, E, act, Bc, 1, Ti, Ti, 1, 0))) return rc;
    if ((rc = launch_gemm(h, s, P + h->nnp_e.w, W(h, h->emb_e.t[1]), h->RESE, E, E, ACT_LINEAR, Bc, 1, 1, Ti, 1, 0))) return rc;
  }
  if (c.act) {
    if ((rc = prep(a, nullptr, 0, M, 1, h->nnp_a.Kp, 2, Bc * To))) return rc;
    if ((rc = launch_gemm(h, {{h->HB, h->nnp_a.Kp, h->nnp_a.Kp, 0, 1, 0}}, P + h->nnp_a.w, W(h, h->emb_a.t[1]), h->AEMB, E2, E2, act,
                          Bc, 1, To, To, 1, 0))) return rc;
  }
  Stream2 st;
  st.x = {{h->XB, E, E, 0, 1, 0}};
  st.e = {{h->EB, E, E, 0, 1, 0}};
  // spatial block: x,e = split(Dense(2E)(concat[x,e])) (emulator.py:231-233); the halves stay views of one 2E-wide row
  auto spatial = [&](int blk, int T) -> int {
    for (int l = 0; l < c.n_sp_layer; ++l) {
      const LayerW& w = h->nn_sp[blk][l];
      std::vector<SegH> segs = st.x;
      segs.insert(segs.end(), st.e.begin(), st.e.end());
      float* z = (l & 1) ? h->EA : h->XA;
      if (h->tape.on && !(h->tape.sp[blk][l].xo = z = h->tape.alloc((size_t)Bc * T * 2 * E))) return fail(h, GNNUDS_E_NOMEM, "activation tape exhausted (NN-cat spatial layer)");
      int r = launch_gemm(h, segs, W(h, w.t[0]), W(h, w.t[1]), z, 2 * E, 2 * E, act, Bc, 1, T, T, 1, 0);
      if (r) return r;
      st.x = {{z, 2 * E, E, 0, 1, 0}};
      st.e = {{z + E, 2 * E, E, 0, 1, 0}};
    }
    return 0;
  };
  if ((rc = spatial(0, Ti))) return rc;
  float *xt = nullptr, *et = nullptr;
  if ((rc = run_temporal(h, h->tpx[0], st.x[0], E, Bc, 1, Ti, To, h->TX0, h->TX1, &xt, h->tape.tp[0][0]))) return rc;
  if ((rc = run_temporal(h, h->tpe[0], st.e[0], E, Bc, 1, Ti, To, h->TE0, h->TE1, &et, h->tape.tp[0][1]))) return rc;
  st.x = {{xt, H, H, 0, 1, 0}, {h->BEMB, E2, E2, 0, 1, 0}};
  st.e = {{et, H, H, 0, 1, 0}};
  if (c.act) st.e.push_back({h->AEMB, E2, E2, 0, 1, 0});
  if ((rc = spatial(1, To))) return rc;
  if ((rc = run_temporal(h, h->tpx[1], st.x[0], E, Bc, 1, To, To, h->TX0, h->TX1, &xt, h->tape.tp[1][0]))) return rc;
  if ((rc = run_temporal(h, h->tpe[1], st.e[0], E, Bc, 1, To, To, h->TE0, h->TE1, &et, h->tape.tp[1][1]))) return rc;
  // ---- resnet + heads (emulator.py:312-338) ----
  const int ract = c.resnet ? ACT_LINEAR : act;
  float *XBo = h->XB, *EBo = h->EB;            // reverse pass: the embeddings in XB / EB stay alive, the resnet output goes on the tape
  if (h->tape.on) {
    h->tape.resx = XBo = h->tape.alloc((size_t)Bc * To * E); h->tape.rese = EBo = h->tape.alloc((size_t)Bc * To * E);
    if (!XBo || !EBo) return fail(h, GNNUDS_E_NOMEM, "activation tape exhausted (NN-cat resnet)");
  }
  if ((rc = launch_gemm(h, {{xt, H, H, 0, 1, 0}}, W(h, h->resx.t[0]), W(h, h->resx.t[1]), XBo, E, E, ract, Bc, 1, To, To, 1, 0))) return rc;
  if ((rc = launch_gemm(h, {{et, H, H, 0, 1, 0}}, W(h, h->rese.t[0]), W(h, h->rese.t[1]), EBo, E, E, ract, Bc, 1, To, To, 1, 0))) return rc;
  if (c.resnet) {
    cumsum_res_act_kernel<float><<<grid_for(h, Bc * (E / 4), 256, 16), 256, 0, h->stream>>>(XBo, h->RES, Bc, To, 1, E, act);
    cumsum_res_act_kernel<float><<<grid_for(h, Bc * (E / 4), 256, 16), 256, 0, h->stream>>>(EBo, h->RESE, Bc, To, 1, E, act);
    h->launches += 2;
    CU(cudaGetLastError());
  }
  const long long So = Bc * To;
  const int ldn = h->ldn();
  if ((rc = launch_gemm(h, {{XBo, E, E, 0, 1, 0}}, P + h->nnp_out.w, P + h->nnp_out.b, h->XA, h->nnp_out.Nop, h->nnp_out.Nop,
                        ACT_HARD_SIGMOID, Bc, 1, To, To, 1, 0))) return rc;
  if ((rc = unflat(h->XA, h->nnp_out.Nop, h->OUTN, ldn, 0, So, N, c.n_out))) return rc;
  if (c.if_flood) {
    std::vector<SegH> cur = {{XBo, E, E, 0, 1, 0}};
    float* bufs[2] = {h->EX, h->NX};
    for (int i = 0; i < c.if_flood; ++i) {
      float* fo = bufs[i & 1];
      if (h->tape.on && (i >= 8 || !(h->tape.flood[i] = fo = h->tape.alloc((size_t)So * E2)))) return fail(h, GNNUDS_E_NOMEM, "activation tape exhausted (NN-cat flood chain)");
      if ((rc = launch_gemm(h, cur, W(h, h->flood_chain[i].t[0]), W(h, h->flood_chain[i].t[1]), fo, E2, E2, act, Bc, 1,
                            To, To, 1, 0))) return rc;
      cur = {{fo, E2, E2, 0, 1, 0}};
    }
    if ((rc = launch_gemm(h, cur, P + h->nnp_flood.w, P + h->nnp_flood.b, h->XA, h->nnp_flood.Nop, h->nnp_flood.Nop, ACT_SIGMOID,
                          Bc, 1, To, To, 1, 0))) return rc;
    if ((rc = unflat(h->XA, h->nnp_flood.Nop, h->OUTN, ldn, c.n_out, So, N, 1))) return rc;
  }
  if ((rc = launch_gemm(h, {{EBo, E, E, 0, 1, 0}}, P + h->nnp_eout.w, P + h->nnp_eout.b, h->EA, h->nnp_eout.Nop, h->nnp_eout.Nop,
                        ACT_TANH, Bc, 1, To, To, 1, 0))) return rc;
  if ((rc = unflat(h->EA, h->nnp_eout.Nop, h->OUTE, c.e_out, 0, So, M, c.e_out))) return rc;
  return run_postproc(h, x, b, a, Bc, y, ey, normalised_io);
}

// =========================================================================================
// Tensor-core path (precision = GNNUDS_PREC_BF16_TC): bf16 activations in HBM, every dense
// contraction on tcgen05 (csrc/tc_gemm.cuh), the GCN/GAT aggregation fused into the epilogue
// of the conv GEMM when a sample's rows fit one CTA tile.
// =========================================================================================
struct AggArgs { int kind = 0; int R = 0; int nnz = 0; const int* rp = nullptr; const int* ci = nullptr; const float* val = nullptr;
                 const float* a_self = nullptr; const float* a_neigh = nullptr; };

// cuTensorMapEncodeTiled through the runtime's driver entry point (no -lcuda at link time)
typedef CUresult (*tmap_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
tmap_encode_fn tmap_encoder() {
  static tmap_encode_fn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<tmap_encode_fn>(f);
  }
  return fn;
}

// {8 elements, rows, width/8 panels} view of a [rows][ld] bf16 buffer, box {8, boxrows, kc/8}
bool make_rows_tmap(CUtensorMap* m, const void* base, long long rows, int ld, int width, int boxrows, int kc) {
  tmap_encode_fn enc = tmap_encoder();
  if (!enc) return false;
  cuuint64_t dims[3] = {8, (cuuint64_t)rows, (cuuint64_t)(width / 8)};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 2, 16};
  cuuint32_t box[3] = {8, (cuuint32_t)boxrows, (cuuint32_t)(kc / 8)};
  cuuint32_t es[3] = {1, 1, 1};
  return enc(m, GU_TMAP_16, 3, const_cast<void*>(base), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// SWIZZLE_128B maps for K-major operand tiles: {C, rows} box {64, boxrows}, or {C, rows of a scenario, Bc} box {64, boxrows, 1}
bool make_rows_tmap_sw(CUtensorMap* m, const void* base, long long rows, int ld, int width, int boxrows, long long Bc = 0) {
  tmap_encode_fn enc = tmap_encoder();
  if (!enc) return false;
  cuuint64_t dims[3] = {(cuuint64_t)width, (cuuint64_t)rows, (cuuint64_t)std::max<long long>(Bc, 1)};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)rows * ld * 2};
  cuuint32_t box[3] = {64, (cuuint32_t)boxrows, 1};
  cuuint32_t es[3] = {1, 1, 1};
  return enc(m, GU_TMAP_16, Bc > 0 ? 3 : 2, const_cast<void*>(base), dims, strides, box, es,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// rank-4 map over (Bc, Ts, Nn, C) activations: {8, Ts*Nn rows of a scenario, Bc, C/8}, box {8, 128, 1, kc/8}
bool make_rows_tmap4(CUtensorMap* m, const void* base, long long Bc, long long rows_per_b, int ld, int width, int kc) {
  tmap_encode_fn enc = tmap_encoder();
  if (!enc) return false;
  cuuint64_t dims[4] = {8, (cuuint64_t)rows_per_b, (cuuint64_t)Bc, (cuuint64_t)(width / 8)};
  cuuint64_t strides[3] = {(cuuint64_t)ld * 2, (cuuint64_t)rows_per_b * ld * 2, 16};
  cuuint32_t box[4] = {8, 128, 1, (cuuint32_t)(kc / 8)};
  cuuint32_t es[4] = {1, 1, 1, 1};
  return enc(m, GU_TMAP_16, 4, const_cast<void*>(base), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

uint32_t tc_agg_scratch_bytes(int MT, int Nout, int g, int R, int nnz, bool gat) {
  return (uint32_t)(5 * 128 * MT + 3 * Nout + (gat ? g * nnz : nnz) + (R + 1) + nnz + 8) * 4u;
}

struct HeadArgs { const float* w; const float* b; float* out; int ldo, col, act; };

int launch_tc_gemm(gnnuds_handle* h, const std::vector<SegH>& segs, int w_tensor, const float* bias, void* out, int ldo,
                   int Nout, int act, long long Bc, int Nn, int Td, int Ts, int pitch_d, int off_d,
                   const AggArgs* agg = nullptr, const HeadArgs* head = nullptr) {
  if (segs.empty() || segs.size() > 4) return fail(h, GNNUDS_E_INVALID, "tc_gemm: 1..4 segments");
  if (w_tensor < 0 || w_tensor >= (int)h->img_off.size() || h->img_off[w_tensor] < 0)
    return fail(h, GNNUDS_E_STATE, "tc_gemm: no bf16 image for weight tensor %d", w_tensor);
  TcGemmP p{};
  p.nseg = (int)segs.size();
  int Ktot = 0, kc = 64;
  for (int i = 0; i < p.nseg; ++i) {
    if (segs[i].width % 32) return fail(h, GNNUDS_E_INVALID, "tc_gemm: segment width %d not a multiple of 32", segs[i].width);
    if (segs[i].width % 64) kc = 32;
    if (segs[i].ld % 8) return fail(h, GNNUDS_E_INVALID, "tc_gemm: row pitch %d not a multiple of 8 elements", segs[i].ld);
    p.seg[i] = TcSeg{static_cast<const bf16*>(static_cast<const void*>(segs[i].p)), segs[i].ld, segs[i].width, segs[i].shift,
                     segs[i].pitch, segs[i].off};
    Ktot += segs[i].width;
  }
  if (Nout % 32 || Nout > 256 || Nout < 32) return fail(h, GNNUDS_E_INVALID, "tc_gemm: Nout %d unsupported", Nout);
  if ((long long)Ktot * Nout != h->tens[w_tensor].numel) return fail(h, GNNUDS_E_INVALID, "tc_gemm: weight %d shape mismatch", w_tensor);
  const long long rows = Bc * Td * Nn;
  if (rows == 0) return 0;
  if (rows >= (1ll << 31)) return fail(h, GNNUDS_E_INVALID, "tc_gemm: %lld rows exceed the 32-bit row index", rows);
  if (Nn >= (1 << 20) || Ts >= 4096) return fail(h, GNNUDS_E_INVALID, "tc_gemm: Nn %d / Ts %d exceed the packed row coordinates", Nn, Ts);
  p.Wimg = h->d_wimg + h->img_off[w_tensor];
  p.bias = bias; p.out = static_cast<bf16*>(out); p.ldo = ldo;
  p.Nout = Nout; p.Ktot = Ktot; p.kc = kc; p.act = act;
  p.rows = (unsigned)rows; p.Nn = Nn; p.Td = Td; p.Ts = Ts; p.pitch_d = pitch_d; p.off_d = off_d;
  if (agg && agg->kind) {
    const int R = agg->R;
    const int g = R <= 128 ? 128 / R : 1;
    p.rpt = g * R; p.MT = (p.rpt + 127) / 128;
    p.agg = agg->kind; p.R = R; p.hsb = (Nout + 8) * 2; p.nnz = agg->nnz;
    if (Nout % 64) return fail(h, GNNUDS_E_INVALID, "tc_gemm: aggregation epilogue needs embed_size %% 64 == 0");
    p.row_ptr = agg->rp; p.col_idx = agg->ci; p.val = agg->val; p.a_self = agg->a_self; p.a_neigh = agg->a_neigh;
    if (!bias) return fail(h, GNNUDS_E_INVALID, "tc_gemm: aggregation epilogue needs the conv bias");
    if (Nn != R || pitch_d != Nn || off_d != 0) return fail(h, GNNUDS_E_INVALID, "tc_gemm: aggregation needs sample-contiguous rows");
  } else {
    p.rpt = 128; p.MT = 1;
    if (head) {
      p.head_w = head->w; p.head_b = head->b; p.head_out = head->out; p.head_ldo = head->ldo; p.head_col = head->col;
      p.head_act = head->act;
    }
  }
  if (p.MT * Nout > 256) return fail(h, GNNUDS_E_INVALID, "tc_gemm: accumulator tile %d x %d exceeds TMEM stage", p.MT * 128, Nout);
  p.n_tiles = (unsigned)((rows + p.rpt - 1) / p.rpt);
  p.w_bytes = (uint32_t)Ktot * Nout * 2;
  p.boxrows = std::min(128 * p.MT, (p.rpt + 7) / 8 * 8);
  p.pstrideA = (uint32_t)p.boxrows * 16u;
  p.stage_bytes = 8u * p.pstrideA;
  // time-shifted segments (causal Conv1D): scenario-aligned tiles, every tap a rank-4 TMA box (tc_gemm.cuh)
  bool shifted = false, tma4_ok = h->tma_on && h->conv1d_tma && !(agg && agg->kind) && pitch_d == Nn && off_d == 0 && Ts >= Td;
  for (int i = 0; i < p.nseg; ++i) {
    const SegH& sg = segs[i];
    shifted |= sg.shift != 0;
    tma4_ok &= sg.pitch == Nn && sg.off == 0 && (reinterpret_cast<uintptr_t>(sg.p) & 15) == 0 && sg.shift <= Ts;
  }
  const bool sw_ok = h->sw128 && kc == 64 && p.boxrows % 8 == 0;
  if (shifted && tma4_ok) {
    const long long rows_src = (long long)Ts * Nn, rows_dst = (long long)Td * Nn;
    for (int i = 0; i < p.nseg && tma4_ok; ++i)
      tma4_ok = sw_ok ? make_rows_tmap_sw(&p.tmap[i], segs[i].p, rows_src, segs[i].ld, segs[i].width, 128, Bc)
                      : make_rows_tmap4(&p.tmap[i], segs[i].p, Bc, rows_src, segs[i].ld, segs[i].width, kc);
    if (tma4_ok) {
      p.tpb = (int)((rows_dst + 127) / 128); p.rows_per_b = (unsigned)rows_dst;
      p.n_tiles = (unsigned)(Bc * p.tpb);
      p.sw128 = sw_ok ? 1 : 0;
      for (int i = 0; i < p.nseg; ++i) p.use_tma[i] = 1;
    }
  }
  if (!p.tpb) {
    bool all = h->tma_on && Ts == Td;
    for (int i = 0; i < p.nseg; ++i) {
      const SegH& sg = segs[i];
      all &= sg.shift == 0 && sg.pitch == Nn && sg.off == 0 && (reinterpret_cast<uintptr_t>(sg.p) & 15) == 0;
    }
    if (all && sw_ok) {
      for (int i = 0; i < p.nseg && all; ++i) all = make_rows_tmap_sw(&p.tmap[i], segs[i].p, rows, segs[i].ld, segs[i].width, p.boxrows);
      if (all) { p.sw128 = 1; for (int i = 0; i < p.nseg; ++i) p.use_tma[i] = 1; }
    }
  }
  for (int i = 0; i < p.nseg && !p.tpb && !p.sw128; ++i) {
    const SegH& sg = segs[i];
    p.use_tma[i] = 0;
    if (h->tma_on && sg.shift == 0 && Ts == Td && sg.pitch == Nn && sg.off == 0 && (reinterpret_cast<uintptr_t>(sg.p) & 15) == 0)
      p.use_tma[i] = make_rows_tmap(&p.tmap[i], sg.p, rows, sg.ld, sg.width, p.boxrows, kc) ? 1 : 0;
  }
  p.stage_off = (p.w_bytes + 1023u) & ~1023u;       // SWIZZLE_128B tiles need 1024-byte alignment
  const uint32_t h_bytes = p.agg ? (((uint32_t)p.rpt * p.hsb + 127u) & ~127u) : 0u;
  const uint32_t sc_bytes = p.agg ? tc_agg_scratch_bytes(p.MT, Nout, p.rpt / p.R, p.R, p.nnz, p.agg == 2) : (uint32_t)Nout * 8u + 2048u;
  const uint32_t fixed = p.stage_off + h_bytes + sc_bytes + 256u + 4096u;
  const uint32_t cap = 227u * 1024u;
  if (fixed + 3u * p.stage_bytes > cap) return fail(h, GNNUDS_E_INVALID, "tc_gemm: tile does not fit shared memory");
  p.NS = (int)std::min<uint32_t>(8u, (cap - fixed) / p.stage_bytes);
  p.h_off = p.stage_off + (uint32_t)p.NS * p.stage_bytes + 4096u;
  p.sc_off = p.h_off + h_bytes;
  p.bar_off = (p.sc_off + sc_bytes + 15u) & ~15u;
  const size_t smem = p.bar_off + (size_t)(2 * p.NS + 6) * 8 + 16;
  cudaEvent_t e1 = nullptr;
  { int prc = prof_begin(h, GNNUDS_FAM_GEMM, &e1); if (prc) return prc; }
  if (!h->attr_done[0]) {
    CU(cudaFuncSetAttribute(tc_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    h->attr_done[0] = true;
  }
  const unsigned grid = std::min<unsigned>(p.n_tiles, (unsigned)h->sm_count);
  CU(launch_pdl(h, tc_rows_kernel, dim3(grid), TC_THREADS, smem, p));
  ++h->launches;
  { int prc = prof_end(h, e1); if (prc) return prc; }
  CU(cudaGetLastError());
  return 0;
}

// {C, Nn, Ts, Bc} view of a (Bc, Ts, Nn, C) activation buffer, box {64 columns, G nodes, Ts steps, 1}: the rows of G whole
// sequences arrive interleaved (row = t*G + g) as one SWIZZLE_128B K-major stage (csrc/fused_tp.cuh)
bool make_seq_tmap(CUtensorMap* m, const void* base, long long Bc, int Ts, int Nn, int ld, int width, int G) {
  tmap_encode_fn enc = tmap_encoder();
  if (!enc) return false;
  cuuint64_t dims[4] = {(cuuint64_t)width, (cuuint64_t)Nn, (cuuint64_t)Ts, (cuuint64_t)Bc};
  cuuint64_t strides[3] = {(cuuint64_t)ld * 2, (cuuint64_t)Nn * ld * 2, (cuuint64_t)Ts * Nn * ld * 2};
  cuuint32_t box[4] = {64, (cuuint32_t)G, (cuuint32_t)Ts, 1};
  cuuint32_t es[4] = {1, 1, 1, 1};
  return enc(m, GU_TMAP_16, 4, const_cast<void*>(base), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// ---- fused Conv1D stack (csrc/fused_tp.cuh).  Returns 0 launched, -1 not applicable, > 0 error. ----
int launch_tp_conv(gnnuds_handle* h, const std::vector<LayerW>& L, const SegH& src, int cin, long long Bc, int Nn, int Ts,
                   int Td, float* out, const LayerW* tail = nullptr, int tail_n = 0, int tail_act = 0) {
  const auto& c = h->c;
  const int H = c.hidden, Pc = c.n_tp_layer, k = c.kernel_size, P = Pc + (tail ? 1 : 0);
  if (!h->tp_fused || !h->tma_on || !h->sw128 || c.recurrent != GNNUDS_REC_CONV1D || H != 64 || cin % 64 || cin > 256 || k < 1 || k > 4 ||
      Pc < 2 || P > TP_MAXP || (tail && tail_n != 128) || Ts > 128 || Ts < 1 || Td > Ts || src.shift || src.pitch != Nn || src.off || src.ld % 8 || src.width != cin ||
      (reinterpret_cast<uintptr_t>(src.p) & 15) || Bc <= 0 || Bc * (long long)Nn >= (1ll << 31))
    return -1;
  TpConvP p{};
  p.P = P; p.Pc = Pc; p.k = k; p.cin = cin; p.H = H; p.Ts = Ts; p.Td = Td; p.Nn = Nn; p.act = act_code(c);
  p.Nlast = tail ? tail_n : H; p.act_last = tail ? tail_act : act_code(c);
  p.G = std::max(1, std::min(std::min(128 / Ts, Nn), 256));
  p.tpb = (Nn + p.G - 1) / p.G;
  p.n_tiles = (unsigned)(Bc * p.tpb);
  p.interleave = P <= 4 ? 1 : 0;        // every layer's output in its own panel buffer: layer 1 of tile i+1 runs ahead
  // zero rows in front of the panels: the largest shift a later layer really uses (shifts >= Ts are skipped)
  int maxshift = 0;
  for (int l = 1; l < Pc; ++l) for (int j = 0; j < k; ++j) { const int sft = (k - 1 - j) << l; if (sft < Ts) maxshift = std::max(maxshift, sft); }
  p.pad_rows = (uint32_t)(maxshift * p.G);
  p.pstrideP = (p.pad_rows + 128u) * 16u;
  auto al = [](uint32_t v, uint32_t a) { return (v + a - 1) / a * a; };
  uint32_t off = 0;
  for (int l = 0; l < P; ++l) {
    const LayerW& lw = l < Pc ? L[l] : *tail;
    const int t = lw.t[0];
    const int K = l < Pc ? (l == 0 ? cin : H) * k : H, Nl = l < Pc ? H : tail_n;
    if (t < 0 || t >= (int)h->img_off.size() || h->img_off[t] < 0 || h->tens[t].numel != (int64_t)K * Nl) return -1;
    p.Wimg[l] = h->d_wimg + h->img_off[t]; p.w_bytes[l] = (uint32_t)K * Nl * 2u; p.w_off[l] = off;
    p.bias[l] = W(h, lw.t[1]);
    off = al(off + p.w_bytes[l], 128);
  }
  const uint32_t pan_bytes = (uint32_t)(H / 8) * p.pstrideP;
  const int npan = p.interleave ? P - 1 : std::min(P - 1, 2);
  off = al(off, 128);
  for (int b = 0; b < TP_MAXP; ++b) p.off_pan[b] = off;
  for (int b = 0; b < npan; ++b) { p.off_pan[b] = off; off = al(off + pan_bytes, 128); }
  p.off_bias = off;
  p.off_bar = al(p.off_bias + (uint32_t)P * 128u * 4u, 16);
  const uint32_t bar_bytes = (uint32_t)(2 * 8 + 2 * TP_MAXP + 2) * 8u + 16u;
  p.off_stage = al(p.off_bar + bar_bytes, 1024);
  const uint32_t cap = 227u * 1024u;
  if (p.off_stage + 2u * 16384u > cap) return -1;
  p.NS = (int)std::min<uint32_t>(8u, (cap - p.off_stage) / 16384u);
  // barrier block was sized for 8 stages; the barrier pointers are laid out from NS
  const size_t smem = p.off_stage + (size_t)p.NS * 16384u;
  if ((P - 1) * H + p.Nlast > 512) return -1;
  if (!make_seq_tmap(&p.tm_in, src.p, Bc, Ts, Nn, src.ld, cin, p.G)) return -1;
  p.out = reinterpret_cast<bf16*>(out);
  if (!h->attr_done[6]) {
    CU(cudaFuncSetAttribute(tp_conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    h->attr_done[6] = true;
  }
  cudaEvent_t e1 = nullptr;
  { int prc = prof_begin(h, GNNUDS_FAM_GEMM, &e1); if (prc) return prc; }
  CU(launch_pdl(h, tp_conv_kernel, dim3(std::min<unsigned>(p.n_tiles, (unsigned)h->sm_count)), TP_THREADS, smem, p));
  ++h->launches;
  { int prc = prof_end(h, e1); if (prc) return prc; }
  CU(cudaGetLastError());
  return 0;
}

int run_temporal_tc(gnnuds_handle* h, const std::vector<LayerW>& L, const SegH& src, int cin, long long Bc, int Nn,
                    int Ts, int Td, float* t0, float* t1, float** dst) {
  const auto& c = h->c;
  const int H = c.hidden, P = c.n_tp_layer;
  if (c.recurrent == GNNUDS_REC_CONV1D) {      // the whole Conv1D stack in one kernel when it applies
    const int frc = launch_tp_conv(h, L, src, cin, Bc, Nn, Ts, Td, t0);
    if (frc > 0) return frc;
    if (frc == 0) { *dst = t0; return 0; }
  }
  const float* cur = src.p;
  int cw = cin, cpitch = src.pitch, coff = src.off;
  float* bufs[2] = {t0, t1};
  for (int i = 0; i < P; ++i) {
    const bool last = (i == P - 1);
    const int Tdst = last ? Td : Ts;
    float* o = bufs[i & 1];
    int rc;
    if (c.recurrent == GNNUDS_REC_CONV1D) {
      std::vector<SegH> segs;
      const int k = c.kernel_size, d = 1 << i;
      for (int j = 0; j < k; ++j) segs.push_back({cur, cw, cw, (k - 1 - j) * d, cpitch, coff});
      if ((rc = launch_tc_gemm(h, segs, L[i].t[0], W(h, L[i].t[1]), o, H, H, act_code(c), Bc, Nn, Tdst, Ts, Nn, 0))) return rc;
    } else {
      const bool gru = c.recurrent == GNNUDS_REC_GRU;
      const int G = (gru ? 3 : 4) * H;
      if ((rc = launch_tc_gemm(h, {{cur, cw, cw, 0, cpitch, coff}}, L[i].t[0], W(h, L[i].t[2]), h->HB, G, G, ACT_LINEAR, Bc, Nn, Ts, Ts,
                               Nn, 0))) return rc;
      const long long nseq = Bc * Nn;
      const int g = (int)((nseq + 31) / 32);
      if (gru && h->gru_tc && H == 64 && h->img_off[L[i].t[1]] >= 0 && nseq < (1ll << 31) * 64) {
        // recurrence on tensor cores (csrc/gru_tc.cuh): 128 sequences per CTA tile, state in registers
        GruTcP p{reinterpret_cast<const bf16*>(h->HB), h->d_wimg + h->img_off[L[i].t[1]], W(h, L[i].t[2]) + G,
                 reinterpret_cast<bf16*>(o), nseq, Nn, Ts, Tdst, H, (unsigned)((nseq + 127) / 128)};
        const size_t smem = 2 * (size_t)H * G * 2 + 2 * 128 * H * 2 + G * 4 + 64;
        if (!h->attr_done[1]) { CU(cudaFuncSetAttribute(gru_tc_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); h->attr_done[1] = true; }
        cudaEvent_t e1 = nullptr;
        { int prc = prof_begin(h, GNNUDS_FAM_RECURRENT, &e1); if (prc) return prc; }
        CU(launch_pdl(h, gru_tc_kernel<64>, dim3(std::min<unsigned>(p.n_tiles, (unsigned)h->sm_count)), GRU_THREADS, smem, p));
        { int prc = prof_end(h, e1); if (prc) return prc; }
      } else if (gru) {
        GruP p{h->HB, W(h, L[i].t[1]), W(h, L[i].t[2]) + G, o, nseq, Nn, Ts, Tdst, H};
        const size_t smem = ((size_t)H * G + 32 * H + 32 * G) * sizeof(float);
        CU(cudaFuncSetAttribute(gru_kernel<bf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        cudaEvent_t e1 = nullptr;
        { int prc = prof_begin(h, GNNUDS_FAM_RECURRENT, &e1); if (prc) return prc; }
        gru_kernel<bf16><<<g, 256, smem, h->stream>>>(p);
        { int prc = prof_end(h, e1); if (prc) return prc; }
      } else if (h->gru_tc && H == 64 && h->img_off[L[i].t[1]] >= 0) {
        // LSTM recurrence on tensor cores (csrc/gru_tc.cuh, lstm_tc_kernel): N = 4H accumulator columns
        LstmTcP p{reinterpret_cast<const bf16*>(h->HB), h->d_wimg + h->img_off[L[i].t[1]], reinterpret_cast<bf16*>(o), nseq, Nn, Ts,
                  Tdst, H, (unsigned)((nseq + 127) / 128)};
        const size_t smem = 2 * (size_t)H * G * 2 + 2 * 128 * H * 2 + 64;
        if (!h->attr_done[2]) { CU(cudaFuncSetAttribute(lstm_tc_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); h->attr_done[2] = true; }
        cudaEvent_t e1 = nullptr;
        { int prc = prof_begin(h, GNNUDS_FAM_RECURRENT, &e1); if (prc) return prc; }
        CU(launch_pdl(h, lstm_tc_kernel<64>, dim3(std::min<unsigned>(p.n_tiles, (unsigned)h->sm_count)), GRU_THREADS, smem, p));
        { int prc = prof_end(h, e1); if (prc) return prc; }
      } else {
        LstmP p{h->HB, W(h, L[i].t[1]), o, nseq, Nn, Ts, Tdst, H};
        const size_t smem = ((size_t)H * G + 64 * H + 32 * G) * sizeof(float);
        CU(cudaFuncSetAttribute(lstm_kernel<bf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        cudaEvent_t e1 = nullptr;
        { int prc = prof_begin(h, GNNUDS_FAM_RECURRENT, &e1); if (prc) return prc; }
        lstm_kernel<bf16><<<g, 256, smem, h->stream>>>(p);
        { int prc = prof_end(h, e1); if (prc) return prc; }
      }
      ++h->launches;
      CU(cudaGetLastError());
    }
    cur = o; cw = H; cpitch = Nn; coff = 0;
  }
  *dst = const_cast<float*>(cur);
  return 0;
}

// conv (GCN/GAT) over `segs`: linear part on tensor cores, aggregation fused when the sample fits a tile
int run_conv_tc(gnnuds_handle* h, const std::vector<SegH>& segs, const LayerW& w, float* out, const int* rp, const int* ci,
                const float* val, int nnz, long long Bc, int T, int R, int act) {
  const auto& c = h->c;
  const int E = c.embed;
  const bool gat = c.conv == GNNUDS_CONV_GAT;
  const float* bias = W(h, w.t[gat ? 3 : 1]);
  const int g = R <= 128 ? 128 / R : 1;
  const int MT = (g * R + 127) / 128;
  int Ktot = 0;
  for (const auto& sg : segs) Ktot += sg.width;
  const size_t need = (((size_t)Ktot * E * 2 + 127) & ~(size_t)127) + (((size_t)g * R * (E + 8) * 2 + 127) & ~(size_t)127) +
                      tc_agg_scratch_bytes(MT, E, g, R, nnz, gat) + 256 + 4096 + 3 * 8 * (size_t)std::min(128 * MT, (g * R + 7) / 8 * 8) * 16;
  if (MT * E <= 256 && E % 64 == 0 && need <= 227 * 1024) {
    AggArgs a;
    a.kind = gat ? 2 : 1; a.R = R; a.nnz = nnz; a.rp = rp; a.ci = ci; a.val = val;
    a.a_self = gat ? W(h, w.t[1]) : nullptr; a.a_neigh = gat ? W(h, w.t[2]) : nullptr;
    return launch_tc_gemm(h, segs, w.t[0], bias, out, E, E, act, Bc, R, T, T, R, 0, &a);
  }
  int rc;
  if ((rc = launch_tc_gemm(h, segs, w.t[0], nullptr, h->HB, E, E, ACT_LINEAR, Bc, R, T, T, R, 0))) return rc;
  return launch_agg<bf16>(h, h->HB, out, rp, ci, val, w, Bc * T, R, act);
}

// ---- fused spatial layer, one stream side (csrc/fused_sp.cuh) -----------------------------------
// Returns 0 launched, -1 not applicable (caller falls back to the layer-by-layer kernels), > 0 error.
int launch_sp_conv(gnnuds_handle* h, const std::vector<SegH>& segs, const void* other_dense, int side, int li,
                   const LayerW& w, void* out, const int* rp, const int* ci, const float* val, int nnz, long long S,
                   int R, int Ro, int act) {
  const auto& c = h->c;
  const int E = c.embed, F = E / 2;
  const bool gat = c.conv == GNNUDS_CONV_GAT;
  // sample groups: sp_g consecutive samples form one block-diagonal super-sample; the tensor maps keep the true row
  // counts, so the rows of a partial last group read as zeros and are not stored (rows_total)
  const long long rows_total = S * R, rows_other = S * Ro;
  { const int g = h->sp_g; S = (S + g - 1) / g; R *= g; Ro *= g; }
  if (h->fused_mode != 1 || segs.empty() || segs.size() > 2 || R > 256 || Ro > 256 || R < 1 || F % 32 || E % 64 || E > 128) return -1;
  if (S <= 0 || S * (long long)std::max(R, Ro) >= (1ll << 31)) return -1;
  SpConvP p{};
  p.nseg = (int)segs.size();
  p.Kself = 0;
  for (int i = 0; i < p.nseg; ++i) {
    const SegH& sg = segs[i];
    if (sg.shift || sg.pitch * h->sp_g != R || sg.off || sg.width % 16 || sg.ld % 8 || (reinterpret_cast<uintptr_t>(sg.p) & 15)) return -1;
    p.segw[i] = sg.width; p.Kself += sg.width;
  }
  const int Ktot = p.Kself + F;
  if ((long long)Ktot * E != h->tens[w.t[0]].numel || h->img_off[w.t[0]] < 0) return -1;
  p.S = (int)S; p.R = R; p.Ro = Ro; p.F = F; p.Nout = E; p.act = act; p.gat = gat ? 1 : 0;
  p.Rbox = (R + 7) / 8 * 8; p.Kp = h->ne_kp[side]; p.MT = (R + 127) / 128;
  p.pstrideA = (uint32_t)p.Rbox * 16u; p.pstrideM = h->ne_pstride[side]; p.pstrideX = (uint32_t)p.Kp * 16u;
  p.hsb = (uint32_t)E * 2u;           // swizzled rows, no padding: the h tile fits the bytes of the self tiles
  p.Wimg = h->d_wimg + h->img_off[w.t[0]]; p.w_bytes = (uint32_t)Ktot * E * 2u;
  p.Mimg = h->d_ne_img[side][li]; p.m_bytes = h->ne_img_bytes[side];
  p.bias = W(h, w.t[gat ? 3 : 1]);
  p.a_self = gat ? W(h, w.t[1]) : nullptr; p.a_neigh = gat ? W(h, w.t[2]) : nullptr;
  // the kernel walks the padded CSR (16-byte index / weight loads); `val` non-null only says "GCN"
  p.row_ptr = h->d_prp[side]; p.col_idx = h->d_pci[side]; p.val = val ? h->d_pval[side] : nullptr; p.nnz = nnz = h->pnnz[side];
  p.bins = h->d_bins[side];
  p.out = static_cast<bf16*>(out);
  auto al = [](uint32_t v, uint32_t a) { return (v + a - 1) / a * a; };
  p.off_w = 0;
  p.off_m = al(p.w_bytes, 128);
  p.off_in = al(p.off_m + p.m_bytes, 1024);
  p.self_bytes = (uint32_t)(p.Kself / 8) * p.pstrideA;
  p.sw_x = (h->sw128 && F == 64 && p.Kp % 8 == 0) ? 1 : 0;
  p.xe_off = al(std::max(p.self_bytes, (uint32_t)R * p.hsb), p.sw_x ? 1024 : 128);   // h tile stays clear of the other_dense tile
  p.xe_bytes = (uint32_t)(F / 8) * p.pstrideX;
  p.in_stride = al(std::max(p.xe_off + p.xe_bytes, (uint32_t)R * p.hsb), 1024);
  p.off_nx = p.off_in + 2 * p.in_stride;
  const uint32_t nx_bytes = (uint32_t)(F / 8) * p.pstrideA;
  p.off_sc = al(p.off_nx + nx_bytes, 128);
  const uint32_t RP = 128u * p.MT;
  const uint32_t sc_bytes = (3u * E + 6u * RP + 2u * (uint32_t)nnz) * 4u + ((uint32_t)R + 1u + (uint32_t)nnz) * 4u +
                            ((uint32_t)SP_BINS + 1u + 2u * (uint32_t)R) * 4u + 64u;
  p.off_bar = al(p.off_sc + sc_bytes, 16);
  // operand views of M-tile mt read 128 rows from row mt*128 of every panel: keep those reads inside the allocation
  uint32_t reach = p.off_nx + (uint32_t)(F / 8 - 1) * p.pstrideA + RP * 16u;
  p.sw128 = h->sw128 ? 1 : 0;
  for (int i = 0; i < p.nseg; ++i) if (segs[i].width % 64) p.sw128 = 0;
  p.chunk_bytes = (uint32_t)p.Rbox * 128u;
  // the last SWIZZLE_128B tile of the second buffer is read for 128*MT rows: keep that inside the allocation
  if (p.sw128) reach = std::max(reach, p.off_in + p.in_stride + (uint32_t)(p.Kself / 64 - 1) * p.chunk_bytes + RP * 128u);
  const size_t smem = std::max<size_t>(p.off_bar + 15 * 8 + 16, reach);
  if (smem > 227 * 1024) return -1;
  for (int i = 0; i < p.nseg; ++i) {
    if (p.sw128 ? !make_rows_tmap_sw(&p.tm_self[i], segs[i].p, rows_total, segs[i].ld, segs[i].width, p.Rbox)
                : !make_rows_tmap(&p.tm_self[i], segs[i].p, rows_total, segs[i].ld, segs[i].width, p.Rbox, segs[i].width)) return -1;
  }
  if (p.sw_x ? !make_rows_tmap_sw(&p.tm_other, other_dense, rows_other, F, F, p.Kp)
             : !make_rows_tmap(&p.tm_other, other_dense, rows_other, F, F, p.Kp, F)) return -1;
  p.rows_total = rows_total;
  if (!h->attr_done[3]) {
    CU(cudaFuncSetAttribute(sp_conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    h->attr_done[3] = true;
  }
  const unsigned grid = (unsigned)std::min<long long>(S, h->sm_count);
  // GNNUDS_TRACE_FILE=<path> [GNNUDS_TRACE_LAUNCH=k]: clock64 stamps of CTA 0 of the k-th sp_conv launch
  static int traced = 0;
  const char* tf = getenv("GNNUDS_TRACE_FILE");
  const int want = getenv("GNNUDS_TRACE_LAUNCH") ? atoi(getenv("GNNUDS_TRACE_LAUNCH")) : 0;
  const bool do_trace = tf && S >= 64 * (long long)h->sm_count && traced++ == want;
  p.dbg = getenv("GNNUDS_SP_DBG") ? atoi(getenv("GNNUDS_SP_DBG")) : 0;
  p.max_len = h->sp_max_len[side];
  const size_t tn = 64 * SP_NSTAMP;
  if (do_trace) { CU(cudaMalloc(&p.trace, tn * sizeof(long long))); CU(cudaMemset(p.trace, 0, tn * sizeof(long long))); }
  cudaEvent_t e1 = nullptr;
  { int prc = prof_begin(h, GNNUDS_FAM_SPATIAL, &e1); if (prc) return prc; }
  CU(launch_pdl(h, sp_conv_kernel, dim3(grid), SP_THREADS, smem, p));
  ++h->launches;
  { int prc = prof_end(h, e1); if (prc) return prc; }
  CU(cudaGetLastError());
  if (do_trace) {
    std::vector<long long> t(tn);
    CU(cudaStreamSynchronize(h->stream));
    CU(cudaMemcpy(t.data(), p.trace, tn * sizeof(long long), cudaMemcpyDeviceToHost));
    cudaFree(p.trace);
    if (FILE* f = fopen(tf, "w")) {
      fprintf(f, "# sp_conv R=%d Ro=%d Kself=%d MT=%d: sample, cycles since first stamp: tma_issue in_full ne_issued ne_done nx_ready conv_done h_full graph_start softmax_done gather_done\n", R, Ro, p.Kself, p.MT);
      for (int i = 0; i < 64; ++i) { fprintf(f, "%d", i); for (int k = 0; k < SP_NSTAMP; ++k) fprintf(f, " %lld", t[i * SP_NSTAMP + k] ? t[i * SP_NSTAMP + k] - t[0] : -1); fprintf(f, "\n"); }
      fclose(f);
    }
  }
  return 0;
}

// ---- fused spatial layer, aggregation on tensor cores (csrc/fused_agg.cuh) ----------------------
// Returns 0 launched, -1 not applicable (caller tries sp_conv_kernel next), > 0 error.
int launch_sp_agg(gnnuds_handle* h, const std::vector<SegH>& segs, const void* other_dense, int side, int li,
                  const LayerW& w, void* out, long long S, int R, int Ro, int act) {
  const auto& c = h->c;
  const int E = c.embed, F = E / 2;
  const bool gat = c.conv == GNNUDS_CONV_GAT;
  const long long rows_total = S * R, rows_other = S * Ro;
  { const int g = h->sp_g; S = (S + g - 1) / g; R *= g; Ro *= g; }
  if (h->fused_mode != 1 || !h->sp_agg || !h->sw128 || segs.empty() || segs.size() > 2 || R > 256 || Ro > 256 || R < 1 || E != 128) return -1;
  if (S <= 0 || S * (long long)std::max(R, Ro) >= (1ll << 31) || !h->annz[side] || h->sp_max_len[side] > 20) return -1;
  if (gat ? (int)h->d_cv_img[side].size() <= li : !h->d_aval[side]) return -1;
  SpAggP p{};
  p.nseg = (int)segs.size();
  p.Kself = 0;
  for (int i = 0; i < p.nseg; ++i) {
    const SegH& sg = segs[i];
    if (sg.shift || sg.pitch * h->sp_g != R || sg.off || sg.width % 64 || sg.ld % 8 || (reinterpret_cast<uintptr_t>(sg.p) & 15)) return -1;
    p.segw[i] = sg.width; p.Kself += sg.width;
  }
  const int Ktot = p.Kself + F;
  if ((long long)Ktot * E != h->tens[w.t[0]].numel || h->img_off[w.t[0]] < 0) return -1;
  p.S = (int)S; p.R = R; p.Ro = Ro; p.F = F; p.Nout = E; p.act = act; p.gat = gat ? 1 : 0;
  p.Rbox = (R + 7) / 8 * 8; p.Kp = h->ne_kp[side]; p.KpJ = (R + 15) / 16 * 16; p.MT = (R + 127) / 128;
  if (p.Kp % 8) return -1;
  // 128 < R <= 144: transposed kernel (features on M, rows on N), csrc/fused_aggt.cuh
  const bool T = h->sp_aggt && p.MT == 2 && p.KpJ <= 144 && p.Kself <= 128 && (!gat || (int)h->d_wa[side].size() > li);
  p.NW = gat && !T ? E + 16 : E;
  p.nb_self = p.MT > 1 ? 1 : 2; p.alias_acc = p.MT > 1 ? 1 : 0;
  p.max_len = h->sp_max_len[side];
  p.pstrideP = (uint32_t)p.Rbox * 16u; p.pstrideM = h->ne_pstride[side]; p.pstrideH = (uint32_t)p.KpJ * 16u; p.pstrideAl = (uint32_t)p.Rbox * 16u;
  if (gat && !T) { p.Wimg = h->d_cv_img[side][li]; p.w_bytes = h->cv_img_bytes[side][li]; }
  else { p.Wimg = h->d_wimg + h->img_off[w.t[0]]; p.w_bytes = (uint32_t)Ktot * E * 2u; }
  if (p.w_bytes != (uint32_t)(Ktot / 8) * (uint32_t)p.NW * 16u) return -1;
  p.Mimg = h->d_ne_img[side][li]; p.m_bytes = h->ne_img_bytes[side];
  p.bias = W(h, w.t[gat ? 3 : 1]);
  p.rp = h->d_arp[side]; p.cj = h->d_acj[side]; p.val = gat ? nullptr : h->d_aval[side]; p.nnz = h->annz[side];
  p.out = static_cast<bf16*>(out); p.rows_total = rows_total;
  if (!h->d_zeros) { CU(cudaMalloc(&h->d_zeros, 65536)); CU(cudaMemset(h->d_zeros, 0, 65536)); }
  p.zeros = h->d_zeros;
  auto al = [](uint32_t v, uint32_t a) { return (v + a - 1) / a * a; };
  const uint32_t RP = 128u * p.MT;
  p.chunk_bytes = (uint32_t)p.Rbox * 128u;
  p.self_bytes = (uint32_t)(p.Kself / 64) * p.chunk_bytes;
  p.xe_bytes = (uint32_t)p.Kp * 128u;
  p.nx_bytes = (uint32_t)(F / 8) * p.pstrideP;
  p.al_bytes = (uint32_t)(p.KpJ / 8) * p.pstrideAl;
  const uint32_t h_bytes = (uint32_t)((p.KpJ + 63) / 64) * (uint32_t)E * 128u;   // h^T tiles [E x 64 neighbours]
  SpAggTP pt{};
  p.off_w = 0;
  if (T) {
    // W^T lives in tensor memory; h^T and the staging tile of the TMA store have their own regions (fused_aggt.cuh)
    p.off_m = 0;
    p.off_self = al(p.m_bytes, 1024);
    p.self_stride = al(p.self_bytes, 1024);
    pt.off_h = p.off_self + p.self_stride;
    pt.off_stage = pt.off_h + al(2u * (uint32_t)E * 128u + (uint32_t)((p.KpJ - 128) / 8) * (uint32_t)E * 16u, 1024);
    p.off_xe = pt.off_stage + 2u * al((uint32_t)R * 128u, 1024);
  } else {
    p.off_m = al(p.w_bytes, 128);
    p.off_self = al(p.off_m + p.m_bytes, 1024);
    p.self_stride = al(std::max(p.self_bytes, h_bytes), 1024);
    p.off_xe = p.off_self + (uint32_t)p.nb_self * p.self_stride;
  }
  p.off_nx = al(p.off_xe + p.xe_bytes, 128);
  p.off_al = al(p.off_nx + p.nx_bytes, 1024);      // doubles as the SWIZZLE_128B staging tile of the TMA store (not T)
  p.stage_half = al((uint32_t)R * 128u, 1024);
  p.al_zero = T ? al(p.al_bytes, 16) : al(std::max(p.al_bytes, 2u * p.stage_half), 16);   // T: own staging tile
  p.off_sc = al(p.off_al + p.al_zero, 128);
  p.agg_n128 = getenv("GNNUDS_AGG_N128") ? atoi(getenv("GNNUDS_AGG_N128")) : 1;
  const uint32_t sc_bytes = ((uint32_t)E + 3u * RP) * 4u + ((uint32_t)R + 1u + (uint32_t)p.nnz * (gat ? 1u : 2u)) * 4u;
  pt.off_wa = al(p.off_sc + sc_bytes, 128);                                  // T: the two score operand images
  pt.off_dump = pt.off_wa + (T && gat ? (uint32_t)(Ktot / 8) * 128u : 0u);
  p.off_bar = T ? al(pt.off_dump + 256u, 16) : al(p.off_sc + sc_bytes, 16);
  // operand views of an M tile read 128 rows from the tile's first row of every panel / swizzle chunk: keep those
  // reads inside the allocation
  uint32_t reach = p.off_bar + 18u * 8u + 16u;
  reach = std::max(reach, p.off_self + (uint32_t)(T ? 0 : p.nb_self - 1) * p.self_stride + (uint32_t)(p.Kself / 64 - 1) * p.chunk_bytes + RP * 128u);
  reach = std::max(reach, p.off_nx + (uint32_t)(F / 8 - 1) * p.pstrideP + RP * 16u);
  reach = std::max(reach, p.off_al + (uint32_t)(p.KpJ / 8 - 1) * p.pstrideAl + RP * 16u);
  reach = std::max(reach, p.off_m + (uint32_t)(p.Kp / 8 - 1) * p.pstrideM + RP * 16u);
  const size_t smem = reach;
  if (smem > 227 * 1024) return -1;
  for (int i = 0; i < p.nseg; ++i)
    if (!make_rows_tmap_sw(&p.tm_self[i], segs[i].p, rows_total, segs[i].ld, segs[i].width, p.Rbox)) return -1;
  if (!make_rows_tmap_sw(&p.tm_other, other_dense, rows_other, F, F, p.Kp)) return -1;
  if (!make_rows_tmap_sw(&p.tm_out, out, rows_total, E, E, R)) return -1;
  if (!h->attr_done[5]) {
    CU(cudaFuncSetAttribute(sp_agg_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    CU(cudaFuncSetAttribute(sp_aggt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    h->attr_done[5] = true;
  }
  const unsigned grid = (unsigned)std::min<long long>(S, h->sm_count);
  static int traced = 0;
  const char* tf = getenv("GNNUDS_TRACE_FILE");
  const int want = getenv("GNNUDS_TRACE_LAUNCH") ? atoi(getenv("GNNUDS_TRACE_LAUNCH")) : 0;
  const bool do_trace = tf && S >= 64 * (long long)h->sm_count && traced++ == want;
  const size_t tn = 64 * SA_NSTAMP;
  if (do_trace) { CU(cudaMalloc(&p.trace, tn * sizeof(long long))); CU(cudaMemset(p.trace, 0, tn * sizeof(long long))); }
  cudaEvent_t e1 = nullptr;
  { int prc = prof_begin(h, GNNUDS_FAM_SPATIAL, &e1); if (prc) return prc; }
  if (T) {
    pt.b = p; pt.wa = gat ? h->d_wa[side][li] : nullptr; pt.Nr = p.KpJ;
    CU(launch_pdl(h, sp_aggt_kernel, dim3(grid), SA_THREADS, smem, pt));
  } else {
    CU(launch_pdl(h, sp_agg_kernel, dim3(grid), SA_THREADS, smem, p));
  }
  ++h->launches;
  { int prc = prof_end(h, e1); if (prc) return prc; }
  CU(cudaGetLastError());
  if (do_trace) {
    std::vector<long long> t(tn);
    CU(cudaStreamSynchronize(h->stream));
    CU(cudaMemcpy(t.data(), p.trace, tn * sizeof(long long), cudaMemcpyDeviceToHost));
    cudaFree(p.trace);
    if (FILE* f = fopen(tf, "w")) {
      fprintf(f, "# sp_agg%s R=%d Ro=%d Kself=%d MT=%d smem=%zu: sample, cycles since first stamp: 0 conv_done_seen 1 h_ready 2 alpha_ready 3 ne_drained 4 agg_done_seen 5 out_done | 6 A:tmem 7 A:stored 8 B:bar1 9 B:zeroed 10 D:tmem 11 D:staged 12 D:bar | MMA thread: 13 conv issued 14 agg issue 15 agg issued 16 NE issue 17 NE issued 18 conv issue\n", T ? "t" : "", R, Ro, p.Kself, p.MT, smem);
      for (int i = 0; i < 64; ++i) { fprintf(f, "%d", i); for (int k = 0; k < SA_NSTAMP; ++k) fprintf(f, " %lld", t[i * SA_NSTAMP + k] ? t[i * SA_NSTAMP + k] - t[0] : -1); fprintf(f, "\n"); }
      fclose(f);
    }
  }
  return 0;
}


int run_spatial_tc(gnnuds_handle* h, int blk, Stream2& st, long long Bc, int T) {
  const auto& c = h->c;
  const int E = c.embed, E2 = E / 2, N = c.n_node, M = c.n_edge;
  const long long S = Bc * T;
  const int act = act_code(c);
  for (int l = 0; l < c.n_sp_layer; ++l) {
    const SpatialW& w = h->sp[blk][l];
    int rc;
    if ((rc = par2(h, [&] { return launch_tc_gemm(h, st.x, w.ex.t[0], W(h, w.ex.t[1]), h->EX, E2, E2, act, Bc, N, T, T, N, 0); },
                   [&] { return launch_tc_gemm(h, st.e, w.xe.t[0], W(h, w.xe.t[1]), h->XE, E2, E2, act, Bc, M, T, T, M, 0); }))) return rc;
    const int li = blk * c.n_sp_layer + l;
    float* xo = (l & 1) ? h->XB : h->XA;
    float* eo = (l & 1) ? h->EB : h->EA;
    // fused NodeEdge + conv + aggregation per stream side; layer-by-layer kernels when it does not apply
    const bool dyn_x = blk == 1 && h->use_adj;   // per-sample attention mask: layer-by-layer kernels on the node side
    auto node_side = [&]() -> int {
      int r = dyn_x ? -1 : launch_sp_agg(h, st.x, h->XE, 0, li, w.conv_x, xo, S, N, M, act);
      if (r < 0 && !dyn_x) r = launch_sp_conv(h, st.x, h->XE, 0, li, w.conv_x, xo, h->d_rp, h->d_ci, h->d_val, h->nnz, S, N, M, act);
      if (r > 0) return r;
      if (r < 0) {
        if ((r = launch_nodeedge<bf16>(h, h->d_ne_matT[0][li], h->ldm_n, h->ti_n, h->XE, h->NX, S, M, N, E2))) return r;
        std::vector<SegH> sx = st.x; sx.push_back({h->NX, E2, E2, 0, N, 0});
        if (dyn_x) {
          if ((r = launch_tc_gemm(h, sx, w.conv_x.t[0], nullptr, h->HB, E, E, ACT_LINEAR, Bc, N, T, T, N, 0))) return r;
          if ((r = launch_agg<bf16>(h, h->HB, xo, h->d_rp, h->d_ci, h->d_val, w.conv_x, S, N, act, true))) return r;
        } else {
          if ((r = run_conv_tc(h, sx, w.conv_x, xo, h->d_rp, h->d_ci, h->d_val, h->nnz, Bc, T, N, act))) return r;
        }
      }
      return 0;
    };
    auto edge_side = [&]() -> int {
      int r = launch_sp_agg(h, st.e, h->EX, 1, li, w.conv_e, eo, S, M, N, act);
      if (r < 0) r = launch_sp_conv(h, st.e, h->EX, 1, li, w.conv_e, eo, h->d_erp, h->d_eci, h->d_eval, h->ennz, S, M, N, act);
      if (r > 0) return r;
      if (r < 0) {
        if ((r = launch_nodeedge<bf16>(h, h->d_ne_matT[1][li], h->ldm_m, h->ti_m, h->EX, h->NE, S, N, M, E2))) return r;
        std::vector<SegH> se = st.e; se.push_back({h->NE, E2, E2, 0, M, 0});
        if ((r = run_conv_tc(h, se, w.conv_e, eo, h->d_erp, h->d_eci, h->d_eval, h->ennz, Bc, T, M, act))) return r;
      }
      return 0;
    };
    if ((rc = par2(h, node_side, edge_side))) return rc;
    st.x = {{xo, E, E, 0, N, 0}};
    st.e = {{eo, E, E, 0, M, 0}};
    if (h->debug_stop == (blk ? 40 : 20) + l) return 1;
  }
  return 0;
}

int forward_chunk_tc(gnnuds_handle* h, const float* x, const float* b, const float* a, const float* ex, long long Bc,
                     float* y, float* ey, bool normalised_io) {
  const auto& c = h->c;
  const int E = c.embed, E2 = E / 2, H = c.hidden, N = c.n_node, M = c.n_edge, Ti = c.seq_in, To = c.seq_out;
  const int act = act_code(c);
  int rc;
  if ((rc = side_prepare(h, Bc))) return rc;
  if ((rc = par2(h,
      [&]() -> int {
        int r;
        if ((r = launch_embed<bf16>(h, x, normalised_io ? nullptr : h->d_nx, h->cx_n, N, nullptr, 0, h->emb_x, h->XB, h->RES, Bc, Ti, N,
                                    c.n_in, E, act, 0))) return r;
        return launch_embed<bf16>(h, b, normalised_io ? nullptr : h->d_nb, h->cb_n, N, nullptr, 0, h->emb_b, h->BEMB, nullptr, Bc, To, N,
                                  c.b_in, E2, act, 0);
      },
      [&]() -> int {
        int r;
        if ((r = launch_embed<bf16>(h, ex, normalised_io ? nullptr : h->d_ne, h->ce_n, M, nullptr, 0, h->emb_e, h->EB, h->RESE, Bc, Ti, M,
                                    c.e_in, E, act, 0))) return r;
        return c.act ? launch_embed<bf16>(h, a, nullptr, 0, 0, h->d_act_of_edge, c.n_act, h->emb_a, h->AEMB, nullptr, Bc, To, M, 1, E2, act, 2) : 0;
      }))) return rc;
  Stream2 st;
  st.x = {{h->XB, E, E, 0, N, 0}};
  st.e = {{h->EB, E, E, 0, M, 0}};
  set_dbg(h, st, Bc * Ti * N, Bc * Ti * M);
  if (h->debug_stop == 1) return 0;
  rc = c.graph_base ? run_spatial_joint<true>(h, 0, st, Bc, Ti) : run_spatial_tc(h, 0, st, Bc, Ti);
  set_dbg(h, st, Bc * Ti * N, Bc * Ti * M);
  if (rc == 1) return 0;
  if (rc) return rc;
  if (h->debug_stop == 2) return 0;
  float *xt = nullptr, *et = nullptr;
  if ((rc = par2(h, [&] { return run_temporal_tc(h, h->tpx[0], st.x[0], E, Bc, N, Ti, To, h->TX0, h->TX1, &xt); },
                 [&] { return run_temporal_tc(h, h->tpe[0], st.e[0], E, Bc, M, Ti, To, h->TE0, h->TE1, &et); }))) return rc;
  st.x = {{xt, H, H, 0, N, 0}, {h->BEMB, E2, E2, 0, N, 0}};
  st.e = {{et, H, H, 0, M, 0}};
  if (c.act) st.e.push_back({h->AEMB, E2, E2, 0, M, 0});
  set_dbg(h, st, Bc * To * N, Bc * To * M);
  if (h->debug_stop == 3) return 0;
  rc = c.graph_base ? run_spatial_joint<true>(h, 1, st, Bc, To) : run_spatial_tc(h, 1, st, Bc, To);
  set_dbg(h, st, Bc * To * N, Bc * To * M);
  if (rc == 1) return 0;
  if (rc) return rc;
  if (h->debug_stop == 4) return 0;
  const int ract = c.resnet ? ACT_LINEAR : act;
  // Conv1D: the block-2 temporal net AND the Dense that follows it (dense_resx, emulator.py:313 / :317) in one kernel per
  // stream - the (B,To,Nn,H) tensor between them never reaches HBM (not under a debug stop, which wants to see it)
  bool tail_x = false, tail_e = false;
  // x_out / e_out (B,To,Nn,E): the buffer of the pair that the block-2 spatial output does NOT occupy
  float* XO = st.x[0].p == h->XA ? h->XB : h->XA;
  float* EO = st.e[0].p == h->EA ? h->EB : h->EA;
  // (measured at C2: 159.4 against 157.1 ms per step without it - the extra weights leave 3 instead of 5 layer-1 stages
  //  in shared memory - so it is opt-in: GNNUDS_TP_FUSED=2)
  if (h->tp_fused == 2 && c.recurrent == GNNUDS_REC_CONV1D && !h->debug_stop && E == 128 && !c.graph_base) {
    int frc = launch_tp_conv(h, h->tpx[1], st.x[0], E, Bc, N, To, To, XO, &h->resx, E, ract);
    if (frc > 0) return frc;
    tail_x = frc == 0;
    frc = launch_tp_conv(h, h->tpe[1], st.e[0], E, Bc, M, To, To, EO, &h->rese, E, ract);
    if (frc > 0) return frc;
    tail_e = frc == 0;
  }
  if ((rc = par2(h, [&] { return tail_x ? 0 : run_temporal_tc(h, h->tpx[1], st.x[0], E, Bc, N, To, To, h->TX0, h->TX1, &xt); },
                 [&] { return tail_e ? 0 : run_temporal_tc(h, h->tpe[1], st.e[0], E, Bc, M, To, To, h->TE0, h->TE1, &et); }))) return rc;
  if (!tail_x) st.x = {{xt, H, H, 0, N, 0}};
  if (!tail_e) st.e = {{et, H, H, 0, M, 0}};
  if (!tail_x && !tail_e) set_dbg(h, st, Bc * To * N, Bc * To * M);
  if (h->debug_stop == 5) return 0;
  if ((rc = par2(h, [&] { return tail_x ? 0 : launch_tc_gemm(h, st.x, h->resx.t[0], W(h, h->resx.t[1]), XO, E, E, ract, Bc, N, To, To, N, 0); },
                 [&] { return tail_e ? 0 : launch_tc_gemm(h, st.e, h->rese.t[0], W(h, h->rese.t[1]), EO, E, E, ract, Bc, M, To, To, M, 0); }))) return rc;
  const long long RXo = Bc * To * N, REo = Bc * To * M;
  const int ldn = h->ldn();
  // resnet + narrow heads in one pass per stream (E == 128): x = act(cumsum_t(x_out) + res) feeds the hard_sigmoid /
  // tanh head straight from registers; the edge stream's x is not written back at all (nothing else reads it),
  // the node stream's only when the flood chain needs it
  const bool fuse_ch = c.resnet && E == 128 && h->fuse_heads && !h->debug_stop && c.n_out <= 4 && c.e_out <= 4;
  if (fuse_ch) {
    CumHeadP px{reinterpret_cast<bf16*>(XO), h->RES, Bc, To, N, act, c.if_flood ? 1 : 0,
                W(h, h->head_out.t[0]), W(h, h->head_out.t[1]), h->OUTN, ldn, 0, c.n_out, ACT_HARD_SIGMOID};
    CumHeadP pe{reinterpret_cast<bf16*>(EO), h->RESE, Bc, To, M, act, 0,
                W(h, h->head_e.t[0]), W(h, h->head_e.t[1]), h->OUTE, c.e_out, 0, c.e_out, ACT_TANH};
    if ((rc = par2(h,
        [&]() -> int {
          if (c.n_out == 1) cumsum_head_bf16_kernel<1, 8><<<grid_for(h, Bc * N, 8, 8), 256, 0, h->stream>>>(px);
          else cumsum_head_bf16_kernel<4, 4><<<grid_for(h, Bc * N, 8, 8), 256, 0, h->stream>>>(px);
          ++h->launches;
          CU(cudaGetLastError());
          return 0;
        },
        [&]() -> int {
          cumsum_head_bf16_kernel<4, 4><<<grid_for(h, Bc * M, 8, 8), 256, 0, h->stream>>>(pe);
          ++h->launches;
          CU(cudaGetLastError());
          return 0;
        }))) return rc;
  } else if (c.resnet) {
    cumsum_res_act_kernel<bf16><<<grid_for(h, Bc * N * (E / 4), 256, 16), 256, 0, h->stream>>>(reinterpret_cast<bf16*>(XO), h->RES, Bc, To, N, E, act);
    cumsum_res_act_kernel<bf16><<<grid_for(h, Bc * M * (E / 4), 256, 16), 256, 0, h->stream>>>(reinterpret_cast<bf16*>(EO), h->RESE, Bc, To, M, E, act);
    h->launches += 2;
    CU(cudaGetLastError());
  }
  st.x = {{XO, E, E, 0, N, 0}};
  st.e = {{EO, E, E, 0, M, 0}};
  set_dbg(h, st, Bc * To * N, Bc * To * M);
  if (h->debug_stop == 6) return 0;
  if (!fuse_ch && (rc = launch_head<bf16>(h, XO, E, E, h->head_out, h->OUTN, ldn, 0, c.n_out, ACT_HARD_SIGMOID, RXo))) return rc;
  if (c.if_flood) {
    std::vector<SegH> cur = st.x;
    float* bufs[2] = {h->EX, h->NX};
    // the last Dense of the chain carries the 1-column sigmoid head in its epilogue (its E/2-wide rows never reach HBM)
    const HeadArgs fh{W(h, h->head_flood.t[0]), W(h, h->head_flood.t[1]), h->OUTN, ldn, c.n_out, ACT_SIGMOID};
    const bool fuse_head = !h->debug_stop && h->fuse_heads;
    for (int i = 0; i < c.if_flood; ++i) {
      const bool lastd = fuse_head && i == c.if_flood - 1;
      if ((rc = launch_tc_gemm(h, cur, h->flood_chain[i].t[0], W(h, h->flood_chain[i].t[1]), bufs[i & 1], E2, E2, act, Bc, N, To, To,
                               N, 0, nullptr, lastd ? &fh : nullptr))) return rc;
      cur = {{bufs[i & 1], E2, E2, 0, N, 0}};
    }
    if (!fuse_head && (rc = launch_head<bf16>(h, cur[0].p, E2, E2, h->head_flood, h->OUTN, ldn, c.n_out, 1, ACT_SIGMOID, RXo))) return rc;
  }
  if (!fuse_ch && (rc = launch_head<bf16>(h, EO, E, E, h->head_e, h->OUTE, c.e_out, 0, c.e_out, ACT_TANH, REo))) return rc;
  return run_postproc(h, x, b, a, Bc, y, ey, normalised_io);
}


// ---- spatial block, graph_base != 0 (emulator.py:220-223 / 273-276): one conv over the joint
// node+edge graph.  z = Conv([x ; e] rows, Adj (N+M,N+M)); x,e = split(z).  Node rows and edge rows of
// a (b,t) sample sit next to each other in a joint buffer (pitch N+M, edge offset N), so the split is free.
template <bool TC>
int run_spatial_joint(gnnuds_handle* h, int blk, Stream2& st, long long Bc, int T) {
  const auto& c = h->c;
  const int E = c.embed, N = c.n_node, M = c.n_edge, R = N + M;
  const long long S = Bc * T;
  const int act = act_code(c);
  for (int l = 0; l < c.n_sp_layer; ++l) {
    const SpatialW& w = h->sp[blk][l];
    int rc;
    float* jo = (l & 1) ? h->J1 : h->J0;
    float* hbj = h->HBJ;
    if (!TC && h->tape.on) {            // reverse pass: h = [x ; e] @ W and the layer output stay alive on the tape
      SpTape& tp = h->tape.sp[blk][l];
      tp.HBx = hbj = h->tape.alloc((size_t)S * R * E); tp.xo = jo = h->tape.alloc((size_t)S * R * E);
      if (!hbj || !jo) return fail(h, GNNUDS_E_NOMEM, "activation tape exhausted (joint spatial layer)");
    }
    if (TC && h->joint_per_sample && (size_t)R * E * 2 >= (8u << 20) && !h->use_adj) {
      // Large joint graph: ONE (b,t) sample at a time - linear part of the sample into a single-sample h buffer, then its
      // gather.  h (51 MB at C5) is re-read ~4x by the neighbour gather; produced and consumed back to back it is still
      // in the 126 MB L2, whereas the whole-chunk schedule wrote every sample's h to HBM first and gathered it back cold.
      auto at = [&](const std::vector<SegH>& sg, long long smp) {
        std::vector<SegH> o = sg;
        for (SegH& q : o) q.p = reinterpret_cast<const float*>(reinterpret_cast<const bf16*>(q.p) + (size_t)smp * q.pitch * q.ld);
        return o;
      };
      for (long long smp = 0; smp < S; ++smp) {
        if ((rc = launch_tc_gemm(h, at(st.x, smp), w.joint.t[0], nullptr, h->HBJ, E, E, ACT_LINEAR, 1, N, 1, 1, R, 0))) return rc;
        if ((rc = launch_tc_gemm(h, at(st.e, smp), w.joint.t[0], nullptr, h->HBJ, E, E, ACT_LINEAR, 1, M, 1, 1, R, N))) return rc;
        float* jos = reinterpret_cast<float*>(reinterpret_cast<bf16*>(jo) + (size_t)smp * R * E);
        if ((rc = launch_agg<bf16>(h, h->HBJ, jos, h->d_rp, h->d_ci, h->d_val, w.joint, 1, R, act, false))) return rc;
      }
    } else if (TC) {
      if ((rc = launch_tc_gemm(h, st.x, w.joint.t[0], nullptr, h->HBJ, E, E, ACT_LINEAR, Bc, N, T, T, R, 0))) return rc;
      if ((rc = launch_tc_gemm(h, st.e, w.joint.t[0], nullptr, h->HBJ, E, E, ACT_LINEAR, Bc, M, T, T, R, N))) return rc;
      if ((rc = launch_agg<bf16>(h, h->HBJ, jo, h->d_rp, h->d_ci, h->d_val, w.joint, S, R, act, blk == 1))) return rc;
    } else {
      if ((rc = launch_gemm(h, st.x, W(h, w.joint.t[0]), nullptr, hbj, E, E, ACT_LINEAR, Bc, N, T, T, R, 0))) return rc;
      if ((rc = launch_gemm(h, st.e, W(h, w.joint.t[0]), nullptr, hbj, E, E, ACT_LINEAR, Bc, M, T, T, R, N))) return rc;
      if ((rc = launch_agg<float>(h, hbj, jo, h->d_rp, h->d_ci, h->d_val, w.joint, S, R, act, blk == 1))) return rc;
    }
    st.x = {{jo, E, E, 0, R, 0}};
    st.e = {{jo, E, E, 0, R, N}};
    if (h->debug_stop == (blk ? 40 : 20) + l) return 1;
  }
  return 0;
}


int forward_any(gnnuds_handle* h, const float* x, const float* b, const float* a, const float* ex, long long Bc, float* y,
                float* ey, bool normalised_io) {
  h->cur_a = a;                 // use_adj: the settings of this chunk drive the block-2 attention mask
  if (h->nn()) return forward_chunk_nn(h, x, b, a, ex, Bc, y, ey, normalised_io);
  return h->tc() ? forward_chunk_tc(h, x, b, a, ex, Bc, y, ey, normalised_io) : forward_chunk(h, x, b, a, ex, Bc, y, ey, normalised_io);
}

// NN-cat keeps ONE row per (b,t): the activation buffers are sized for nn_rows_equiv() "nodes" of width E, enough for
// the widest flattened row (inputs N*C, 2E spatial rows, flattened heads)
int64_t nn_rows_equiv(const gnnuds_config& c) {
  auto r16 = [](int64_t v) { return (v + 15) / 16 * 16; };
  const int64_t N = c.n_node, M = c.n_edge;
  int64_t w = 2 * c.embed;
  for (int64_t v : {N * c.n_in, N * c.b_in, M * c.e_in, M, N * c.n_out, M * c.e_out, N}) w = std::max(w, r16(v));
  return std::max<int64_t>(2, (w + c.embed - 1) / c.embed);
}

int64_t floats_per_scenario(const gnnuds_config& c) {
  const int64_t E = c.embed, E2 = E / 2, H = c.hidden;
  const bool nn = c.conv == GNNUDS_CONV_NONE;
  int64_t N = nn ? nn_rows_equiv(c) : c.n_node, M = nn ? nn_rows_equiv(c) : c.n_edge;
  const int64_t Tm = std::max(c.seq_in, c.seq_out);
  const int64_t G = c.recurrent == GNNUDS_REC_GRU ? 3 * H : c.recurrent == GNNUDS_REC_LSTM ? 4 * H : 0;
  const int64_t W2 = std::max(E2, H);
  int64_t f = 0;
  f += Tm * (N + M) * (2 * E + 2 * W2 + 2 * H);          // XA XB EA EB | XE EX NX NE | TX* TE*
  f += Tm * std::max(N, M) * std::max(E, G) + Tm * std::max<int64_t>(c.n_node, c.n_edge) * 2;   // HB, SS, SN
  f += c.seq_out * (N + M) * E2;                         // BEMB AEMB
  f += (N + M) * E;                                      // RES RESE
  if (nn) { N = c.n_node; M = c.n_edge; }
  f += c.seq_out * (N + M) * 4;                          // OUTN OUTE
  f += 2 * (c.seq_in * (N * c.n_in + M * c.e_in) + c.seq_out * (N * c.b_in + std::max(c.n_act, 1)));   // two staging sets
  f += 2 * c.seq_out * (N * 5 + M * 4);
  if (c.graph_base) f += Tm * (N + M) * (3 * E + 2);
  return f + 64;
}

int ensure_workspace(gnnuds_handle* h, int64_t Bc) {
  if (Bc <= h->ws_chunk && h->d_ws) return 0;
  const auto& c = h->c;
  ++h->buf_gen;
  if (h->d_ws) { cudaFree(h->d_ws); h->d_ws = nullptr; }
  const int64_t E = c.embed, E2 = E / 2, H = c.hidden, N = c.n_node, M = c.n_edge;
  const int64_t Nw = h->nn() ? nn_rows_equiv(c) : N, Mw = h->nn() ? nn_rows_equiv(c) : M;   // activation rows per (b,t)
  const int64_t Tm = std::max(c.seq_in, c.seq_out);
  const int64_t G = c.recurrent == GNNUDS_REC_GRU ? 3 * H : c.recurrent == GNNUDS_REC_LSTM ? 4 * H : 0;
  const int64_t W2 = std::max(E2, H);
  const int64_t RX = Bc * Tm * Nw, RE = Bc * Tm * Mw, RM = Bc * Tm * std::max(Nw, Mw);
  const int64_t RS = Bc * Tm * std::max(N, M);   // post-processing scratch rows
  const int64_t a = (int64_t)h->es();      // activation element size (fp32 path 4, tensor-core path 2)
  struct Item { float** p; int64_t bytes; };
  std::vector<Item> plan = {
      {&h->XA, RX * E * a}, {&h->XB, RX * E * a}, {&h->EA, RE * E * a}, {&h->EB, RE * E * a}, {&h->HB, RM * std::max(E, G) * a},
      {&h->XE, RE * W2 * a}, {&h->EX, RX * W2 * a}, {&h->NX, RX * W2 * a}, {&h->NE, RE * W2 * a},
      {&h->TX0, RX * H * a}, {&h->TX1, RX * H * a}, {&h->TE0, RE * H * a}, {&h->TE1, RE * H * a},
      {&h->BEMB, Bc * c.seq_out * Nw * E2 * a}, {&h->AEMB, Bc * c.seq_out * Mw * E2 * a},
      {&h->RES, Bc * Nw * E * 4}, {&h->RESE, Bc * Mw * E * 4},
      {&h->OUTN, Bc * c.seq_out * N * 4 * 4}, {&h->OUTE, Bc * c.seq_out * M * 4 * 4}, {&h->SS, RS * 4}, {&h->SN, RS * 4},
      {&h->in_x, Bc * c.seq_in * N * c.n_in * 4}, {&h->in_b, Bc * c.seq_out * N * c.b_in * 4},
      {&h->in_a, Bc * c.seq_out * std::max(c.n_act, 1) * 4}, {&h->in_ex, Bc * c.seq_in * M * c.e_in * 4},
      {&h->out_y, Bc * c.seq_out * N * 5 * 4}, {&h->out_ey, Bc * c.seq_out * M * 4 * 4},
      {&h->in2_x, Bc * c.seq_in * N * c.n_in * 4}, {&h->in2_b, Bc * c.seq_out * N * c.b_in * 4},
      {&h->in2_a, Bc * c.seq_out * std::max(c.n_act, 1) * 4}, {&h->in2_ex, Bc * c.seq_in * M * c.e_in * 4},
      {&h->out2_y, Bc * c.seq_out * N * 5 * 4}, {&h->out2_ey, Bc * c.seq_out * M * 4 * 4}};
  if (c.graph_base) {
    const int64_t RJ = Bc * Tm * (N + M);
    plan.push_back({&h->HBJ, RJ * E * a}); plan.push_back({&h->J0, RJ * E * a}); plan.push_back({&h->J1, RJ * E * a});
    plan[19].bytes = RJ * 4; plan[20].bytes = RJ * 4;   // SS, SN cover the joint rows
  }
  size_t total = 0;
  for (auto& it : plan) total += (size_t)((it.bytes + 255) / 256 * 256);
  if (cudaMalloc(&h->d_ws, total) != cudaSuccess) {
    (void)cudaGetLastError();
    h->d_ws = nullptr; h->ws_chunk = 0;
    return fail(h, GNNUDS_E_NOMEM, "workspace of %zu MB for %lld scenarios per pass does not fit the device (lower max_chunk / workspace_mb)",
                total >> 20, (long long)Bc);
  }
  CU(cudaMemsetAsync(h->d_ws, 0, total, h->stream));
  size_t off = 0;
  for (auto& it : plan) { *it.p = reinterpret_cast<float*>(reinterpret_cast<char*>(h->d_ws) + off); off += (size_t)((it.bytes + 255) / 256 * 256); }
  h->ws_floats = total / 4;
  h->ws_chunk = Bc;
  return 0;
}

int64_t pick_chunk(const gnnuds_handle* h, int64_t B) {
  const auto& c = h->c;
  if (c.max_chunk > 0) return std::min<int64_t>(B, c.max_chunk);
  const int64_t budget = (int64_t)(c.workspace_mb > 0 ? c.workspace_mb : h->default_ws_mb /* include/gnnuds.h: 0 = 8192 MB */) * 1024 * 1024 / 4;
  int64_t bc = budget / floats_per_scenario(c) * (h->tc() ? 2 : 1);
  if (bc < 1) bc = 1;
  return std::min(B, bc);
}

int check_ready(gnnuds_handle* h) {
  if (!h) return GNNUDS_E_INVALID;
  if (!h->has_graph) return fail(h, GNNUDS_E_STATE, "gnnuds_set_graph has not been called");
  if (!h->has_weights) return fail(h, GNNUDS_E_STATE, "gnnuds_set_weights has not been called");
  if (!h->has_norm) return fail(h, GNNUDS_E_STATE, "gnnuds_set_norm has not been called");
  return 0;
}

void begin_timing(gnnuds_handle* h) { h->gemm_ev_used = 0; h->last_total_ms = 0.f; h->last_gemm_ms = 0.f; }

int finish_timing(gnnuds_handle* h) {
  if (!h->profiling) return 0;
  for (int f = 0; f < 4; ++f) { h->fam_ms[f] = 0.f; h->fam_launches[f] = 0; }
  for (size_t i = 0; i < h->gemm_ev_used; ++i) {
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, h->gemm_ev[i].first, h->gemm_ev[i].second));
    h->fam_ms[h->ev_family[i] & 3] += ms;
    ++h->fam_launches[h->ev_family[i] & 3];
  }
  h->last_gemm_ms = h->fam_ms[GNNUDS_FAM_GEMM];
  return 0;
}

template <typename T>
int upload(gnnuds_handle* h, T** dst, const T* src, size_t n) {
  ++h->buf_gen;
  if (*dst) { cudaFree(*dst); *dst = nullptr; }
  if (!n) return 0;
  CU(cudaMalloc(dst, n * sizeof(T)));
  CU(cudaMemcpy(*dst, src, n * sizeof(T), cudaMemcpyHostToDevice));
  return 0;
}


// ---- pageable host buffers ---------------------------------------------------------------------
bool is_pageable(const void* p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { (void)cudaGetLastError(); return true; }
  return at.type == cudaMemoryTypeUnregistered;
}

struct CopyJob { void* dst; const void* src; size_t bytes; };
// memcpy split over a few threads (one core moves ~10 GB/s, less than a step's copies need to hide behind the compute)
void parallel_copy(const std::vector<CopyJob>& jobs) {
  size_t total = 0;
  for (const CopyJob& j : jobs) total += j.bytes;
  const unsigned hw = std::thread::hardware_concurrency();
  int T = (int)std::max(1u, std::min(8u, hw ? hw / 2 : 4u));
  if (total < (8u << 20)) T = 1;
  if (T == 1) {
    for (const CopyJob& j : jobs) if (j.bytes) std::memcpy(j.dst, j.src, j.bytes);
    return;
  }
  std::vector<std::thread> th;
  for (int t = 0; t < T; ++t)
    th.emplace_back([&jobs, t, T]() {
      for (const CopyJob& j : jobs) {
        const size_t n = j.bytes, lo = n * t / T / 64 * 64, hi = t == T - 1 ? n : n * (t + 1) / T / 64 * 64;
        if (hi > lo) std::memcpy(static_cast<char*>(j.dst) + lo, static_cast<const char*>(j.src) + lo, hi - lo);
      }
    });
  for (auto& x : th) x.join();
}

int ensure_pinned(gnnuds_handle* h, size_t in_floats, size_t out_floats) {
  if (in_floats > h->pin_in_floats) {
    for (int k = 0; k < 2; ++k) { if (h->pin_in[k]) cudaFreeHost(h->pin_in[k]); h->pin_in[k] = nullptr; }
    h->pin_in_floats = 0;
    for (int k = 0; k < 2; ++k)
      if (cudaMallocHost(&h->pin_in[k], in_floats * sizeof(float)) != cudaSuccess) { (void)cudaGetLastError(); return fail(h, GNNUDS_E_NOMEM, "pinned staging ring: %zu MB not available", (in_floats * 4) >> 20); }
    h->pin_in_floats = in_floats;
  }
  if (out_floats > h->pin_out_floats) {
    for (int k = 0; k < 2; ++k) { if (h->pin_out[k]) cudaFreeHost(h->pin_out[k]); h->pin_out[k] = nullptr; }
    h->pin_out_floats = 0;
    for (int k = 0; k < 2; ++k)
      if (cudaMallocHost(&h->pin_out[k], out_floats * sizeof(float)) != cudaSuccess) { (void)cudaGetLastError(); return fail(h, GNNUDS_E_NOMEM, "pinned staging ring: %zu MB not available", (out_floats * 4) >> 20); }
    h->pin_out_floats = out_floats;
  }
  return 0;
}


// grow-only device buffer `slot` of the handle with room for n floats
int grow(gnnuds_handle* h, int slot, size_t n, float** out) {
  n = std::max<size_t>(n, 1);
  if (h->gb_cap[slot] < n) {
    if (h->gb[slot]) { cudaStreamSynchronize(h->stream); cudaFree(h->gb[slot]); h->gb[slot] = nullptr; h->gb_cap[slot] = 0; }
    if (cudaMalloc(&h->gb[slot], n * sizeof(float)) != cudaSuccess) { (void)cudaGetLastError(); return fail(h, GNNUDS_E_NOMEM, "device buffer of %zu MB", (n * 4) >> 20); }
    h->gb_cap[slot] = n;
    ++h->buf_gen;
  }
  *out = h->gb[slot];
  return 0;
}

// Decide whether this chunk runs its node and edge sides concurrently (par2) and make the side's scratch exist.
// Called by the entry points BEFORE any graph capture (it may allocate) and again, as a no-op, by the forward itself.
int side_prepare(gnnuds_handle* h, int64_t Bc) {
  const auto& c = h->c;
  const int64_t Tm = std::max(c.seq_in, c.seq_out);
  const int64_t rows = Bc * Tm * std::max(c.n_node, c.n_edge);
  h->side_on = h->side_par && h->s_side && h->profiling < 2 && !h->debug_stop && !h->nn() && !c.graph_base && !h->use_adj &&
               std::max(c.n_node, c.n_edge) <= 256 && rows <= 262144;
  if (!h->side_on) return 0;
  const int64_t G = c.recurrent == GNNUDS_REC_GRU ? 3 * c.hidden : c.recurrent == GNNUDS_REC_LSTM ? 4 * c.hidden : 0;
  float* p = nullptr;
  const int rc = grow(h, 41, (size_t)(rows * std::max<int64_t>(c.embed, G)), &p);     // fp32-sized: covers both precisions
  if (rc) { h->side_on = false; return rc; }
  h->HB2 = p;
  return 0;
}

int ensure_rollout_state(gnnuds_handle* h, int64_t Bc) {
  const auto& c = h->c;
  if (h->ro_B >= Bc && h->ro_x) return 0;
  ++h->buf_gen;
  for (float** p : {&h->ro_x, &h->ro_ex, &h->ro_b, &h->ro_a}) if (*p) { cudaFree(*p); *p = nullptr; }
  h->ro_B = 0;
  const int64_t T = c.seq_out;
  if (cudaMalloc(&h->ro_x, Bc * c.seq_in * c.n_node * c.n_in * sizeof(float)) != cudaSuccess ||
      cudaMalloc(&h->ro_ex, Bc * c.seq_in * c.n_edge * c.e_in * sizeof(float)) != cudaSuccess ||
      cudaMalloc(&h->ro_b, Bc * T * c.n_node * c.b_in * sizeof(float)) != cudaSuccess ||
      cudaMalloc(&h->ro_a, std::max<int64_t>(Bc * T * c.n_act, 1) * sizeof(float)) != cudaSuccess) {
    (void)cudaGetLastError();
    return fail(h, GNNUDS_E_NOMEM, "rollout state buffers");
  }
  h->ro_B = Bc;
  return 0;
}

}  // namespace

// =========================================================================================
// C ABI
// =========================================================================================
extern "C" {

const char* gnnuds_last_error(const gnnuds_handle* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int gnnuds_create(const gnnuds_config* cfg, gnnuds_handle** out) {
  gnnuds_handle* h = nullptr;
  if (!cfg || !out) return fail(nullptr, GNNUDS_E_INVALID, "null argument");
  if (cfg->abi_version != GNNUDS_ABI_VERSION) return fail(nullptr, GNNUDS_E_INVALID, "ABI version mismatch (%d vs %d)", cfg->abi_version, GNNUDS_ABI_VERSION);
  const gnnuds_config& c = *cfg;
  if (c.conv != GNNUDS_CONV_GCN && c.conv != GNNUDS_CONV_GAT && c.conv != GNNUDS_CONV_NONE)
    return fail(nullptr, GNNUDS_E_INVALID, "conv kind %d unsupported (NN-cat/GCN/GAT only; Diffusion/General are not on the CUDA path)", c.conv);
  if (c.conv == GNNUDS_CONV_NONE && (c.n_node < 2 || c.n_edge < 2 || c.graph_base))
    return fail(nullptr, GNNUDS_E_INVALID, "NN-cat needs n_node, n_edge >= 2 and graph_base == 0 (graph_base is ignored without conv, emulator.py:221)");
  if (c.recurrent < GNNUDS_REC_CONV1D || c.recurrent > GNNUDS_REC_LSTM) return fail(nullptr, GNNUDS_E_INVALID, "recurrent kind %d unsupported", c.recurrent);
  if (c.embed % 32 || c.embed < 32 || c.embed > 256) return fail(nullptr, GNNUDS_E_INVALID, "embed_size %d must be a multiple of 32 in [32,256]", c.embed);
  if (c.hidden % 16 || c.hidden > 128 || (c.recurrent != GNNUDS_REC_CONV1D && (c.hidden * 3) % 64)) return fail(nullptr, GNNUDS_E_INVALID, "hidden_dim %d unsupported", c.hidden);
  if (c.hidden % 64) return fail(nullptr, GNNUDS_E_INVALID, "hidden_dim %d must be a multiple of 64", c.hidden);
  if (c.n_in > 8 || c.e_in > 8 || c.b_in > 8 || c.n_in < 1) return fail(nullptr, GNNUDS_E_INVALID, "channel counts out of range");
  if (c.seq_in < c.seq_out) return fail(nullptr, GNNUDS_E_INVALID, "seq_in < seq_out (emulator.py:249 needs seq_in >= seq_out)");
  if (c.act && c.seq_in != c.seq_out) return fail(nullptr, GNNUDS_E_INVALID, "act requires seq_in == seq_out (emulator.py:191)");
  if (c.recurrent == GNNUDS_REC_CONV1D && c.kernel_size > 4) return fail(nullptr, GNNUDS_E_INVALID, "kernel_size > 4 unsupported");
  if (c.n_out + (c.if_flood ? 1 : 0) > 4 || c.e_out > 4) return fail(nullptr, GNNUDS_E_INVALID, "head widths out of range");
  if (c.edge_fusion ? c.n_out != 1 : c.n_out != 3) return fail(nullptr, GNNUDS_E_INVALID, "n_out %d inconsistent with edge_fusion", c.n_out);
  if (c.precision != GNNUDS_PREC_FP32 && c.precision != GNNUDS_PREC_BF16_TC)
    return fail(nullptr, GNNUDS_E_INVALID, "precision %d unknown", c.precision);
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return fail(nullptr, GNNUDS_E_CUDA, "no CUDA device: the gnn_uds_b200 product path has no CPU fallback");
  if (c.device < 0 || c.device >= ndev) return fail(nullptr, GNNUDS_E_INVALID, "device %d out of range (%d devices)", c.device, ndev);
  h = new gnnuds_handle();
  h->c = c;
  // NN-cat is a chain of small row GEMMs over Bc*T rows (no per-node rows): it runs on the fp32 SIMT kernels whatever
  // precision was asked for
  if (c.conv == GNNUDS_CONV_NONE) h->c.precision = GNNUDS_PREC_FP32;
  if (cudaSetDevice(c.device) != cudaSuccess || cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess) {
    g_create_error = "cudaSetDevice/cudaStreamCreate failed";
    delete h;
    return GNNUDS_E_CUDA;
  }
  cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, c.device);
  cudaEventCreate(&h->ev_a); cudaEventCreate(&h->ev_b);
  cudaStreamCreateWithFlags(&h->s_h2d, cudaStreamNonBlocking);
  cudaStreamCreateWithFlags(&h->s_d2h, cudaStreamNonBlocking);
  cudaStreamCreateWithFlags(&h->s_side, cudaStreamNonBlocking);
  cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming);
  cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming);
  for (int i = 0; i < 2; ++i) {
    cudaEventCreateWithFlags(&h->ev_in[i], cudaEventDisableTiming);
    cudaEventCreateWithFlags(&h->ev_comp[i], cudaEventDisableTiming);
    cudaEventCreateWithFlags(&h->ev_out[i], cudaEventDisableTiming);
  }
  h->tma_on = !(getenv("GNNUDS_NO_TMA") && getenv("GNNUDS_NO_TMA")[0] == '1');
  if (getenv("GNNUDS_FUSED")) h->fused_mode = atoi(getenv("GNNUDS_FUSED"));
  if (getenv("GNNUDS_GRU_TC")) h->gru_tc = atoi(getenv("GNNUDS_GRU_TC"));
  if (getenv("GNNUDS_SP_AGG")) h->sp_agg = atoi(getenv("GNNUDS_SP_AGG"));
  if (getenv("GNNUDS_GRAPHS")) h->use_graphs = atoi(getenv("GNNUDS_GRAPHS"));
  if (getenv("GNNUDS_SIDE_PAR")) h->side_par = atoi(getenv("GNNUDS_SIDE_PAR"));
  if (getenv("GNNUDS_SKINNY_ROWS")) h->skinny_rows = atoll(getenv("GNNUDS_SKINNY_ROWS"));
  if (getenv("GNNUDS_PDL")) h->pdl = atoi(getenv("GNNUDS_PDL"));
  if (getenv("GNNUDS_WORKSPACE_MB") && atoi(getenv("GNNUDS_WORKSPACE_MB")) > 0) h->default_ws_mb = atoi(getenv("GNNUDS_WORKSPACE_MB"));
  if (getenv("GNNUDS_JOINT_PER_SAMPLE")) h->joint_per_sample = atoi(getenv("GNNUDS_JOINT_PER_SAMPLE"));
  if (getenv("GNNUDS_TP_FUSED")) h->tp_fused = atoi(getenv("GNNUDS_TP_FUSED"));
  if (getenv("GNNUDS_SP_AGGT")) h->sp_aggt = atoi(getenv("GNNUDS_SP_AGGT"));
  h->sp_g = std::max(1, std::min(8, 128 / std::max(1, std::max(c.n_node, c.n_edge))));
  if (getenv("GNNUDS_SP_GROUP")) h->sp_g = std::max(1, std::min(h->sp_g, atoi(getenv("GNNUDS_SP_GROUP"))));
  if (getenv("GNNUDS_SW128")) h->sw128 = atoi(getenv("GNNUDS_SW128"));
  if (getenv("GNNUDS_CONV1D_TMA")) h->conv1d_tma = atoi(getenv("GNNUDS_CONV1D_TMA"));
  if (getenv("GNNUDS_FUSE_HEADS")) h->fuse_heads = atoi(getenv("GNNUDS_FUSE_HEADS"));
  LayoutBuilder lb{h};
  lb.build();
  *out = h;
  return 0;
}

// Gather schedule of sp_conv_kernel: the R output rows are packed into SP_BINS row lists of nearly equal
// neighbour count (longest-processing-time first), one list per 8-lane group of the graph warps, so that
// a sample's aggregation takes ~nnz/SP_BINS neighbour steps instead of (passes x max row degree).
int upload_bins(gnnuds_handle* h, int side, const int* rp0, const int* ci0, const float* val0, int R0) {
  // block-diagonal replication for the sample groups of small graphs (h->sp_g)
  const int g = h->sp_g, R = g * R0;
  std::vector<int> rpv(R + 1, 0), civ;
  std::vector<float> valv;
  for (int q = 0; q < g; ++q)
    for (int r = 0; r < R0; ++r) {
      for (int k = rp0[r]; k < rp0[r + 1]; ++k) { civ.push_back(ci0[k] + q * R0); if (val0) valv.push_back(val0[k]); }
      rpv[q * R0 + r + 1] = (int)civ.size();
    }
  const int* rp = rpv.data(); const int* ci = civ.data(); const float* val = val0 ? valv.data() : nullptr;
  // padded CSR: every row starts on a multiple of 4 entries; pads repeat the row's first neighbour with weight 0
  {
    std::vector<int> prp(R + 1, 0), pci;
    std::vector<float> pval;
    for (int r = 0; r < R; ++r) {
      const int len = rp[r + 1] - rp[r], plen = (len + 3) / 4 * 4;
      for (int k = 0; k < plen; ++k) {
        pci.push_back(k < len ? ci[rp[r] + k] : (len ? ci[rp[r]] : r));
        pval.push_back(k < len && val ? val[rp[r] + k] : 0.f);
      }
      prp[r + 1] = (int)pci.size();
    }
    if (pci.empty()) { pci.assign(4, 0); pval.assign(4, 0.f); }
    void* old[] = {h->d_prp[side], h->d_pci[side], h->d_pval[side]};
    for (void* q : old) if (q) cudaFree(q);
    h->d_prp[side] = h->d_pci[side] = nullptr; h->d_pval[side] = nullptr;
    int rc;
    if ((rc = upload(h, &h->d_prp[side], prp.data(), prp.size()))) return rc;
    if ((rc = upload(h, &h->d_pci[side], pci.data(), pci.size()))) return rc;
    if ((rc = upload(h, &h->d_pval[side], pval.data(), pval.size()))) return rc;
    h->pnnz[side] = prp[R];
  }
  {  // sp_agg_kernel tables: alpha[i][j] of the dense K-major operand sits at (j/8)*pstrideAl + i*16 + (j%8)*2
    const uint32_t pstrideAl = (uint32_t)((R + 7) / 8 * 8) * 16u;
    std::vector<uint32_t> cj(std::max<size_t>(civ.size(), 1), 0u);
    bool ok = R <= 256;
    for (int r = 0; r < R && ok; ++r)
      for (int k = rp[r]; k < rp[r + 1]; ++k) {
        const uint32_t j = (uint32_t)ci[k], off = (j >> 3) * pstrideAl + (uint32_t)r * 16u + (j & 7u) * 2u;
        if (off >= (1u << 24)) { ok = false; break; }
        cj[k] = (off << 8) | j;
      }
    void* old[] = {h->d_arp[side], h->d_acj[side], h->d_aval[side]};
    for (void* q : old) if (q) cudaFree(q);
    h->d_arp[side] = nullptr; h->d_acj[side] = nullptr; h->d_aval[side] = nullptr; h->annz[side] = 0;
    if (ok) {
      int rc;
      if ((rc = upload(h, &h->d_arp[side], rpv.data(), rpv.size()))) return rc;
      if ((rc = upload(h, &h->d_acj[side], cj.data(), cj.size()))) return rc;
      if (val && !valv.empty()) { if ((rc = upload(h, &h->d_aval[side], valv.data(), valv.size()))) return rc; }
      h->annz[side] = (int)civ.size();
    }
  }
  // The four 8-lane groups of a warp run in lockstep: a warp's time is the sum over its row slots of the LONGEST row
  // of the slot.  So rows are sorted by length, cut into quads of similar length (one quad = one slot of one warp),
  // and the quads are spread over the warps longest-processing-time first.
  std::vector<int> order(R), which(R);
  for (int i = 0; i < R; ++i) order[i] = i;
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return rp[a + 1] - rp[a] > rp[b + 1] - rp[b]; });
  {
    constexpr int NW = SP_BINS / 4;
    std::vector<int> wload(NW, 0);
    for (int q0 = 0; q0 < R; q0 += 4) {
      const int len0 = rp[order[q0] + 1] - rp[order[q0]];
      int best = 0;
      for (int w = 1; w < NW; ++w) if (wload[w] < wload[best]) best = w;
      wload[best] += (len0 + 3) / 4 * 4 + 2;                 // blocks of 4 neighbours + per-row epilogue
      for (int j = 0; j < 4 && q0 + j < R; ++j) which[order[q0 + j]] = best * 4 + j;
    }
  }
  std::vector<int> bins(SP_BINS + 1 + 2 * R, 0);
  for (int r = 0; r < R; ++r) bins[SP_BINS + 1 + R + r] = rp[r + 1] - rp[r];
  h->sp_max_len[side] = 0;
  for (int r = 0; r < R; ++r) h->sp_max_len[side] = std::max(h->sp_max_len[side], rp[r + 1] - rp[r]);
  for (int r = 0; r < R; ++r) ++bins[which[r] + 1];
  for (int b = 0; b < SP_BINS; ++b) bins[b + 1] += bins[b];
  std::vector<int> fill(bins.begin(), bins.begin() + SP_BINS);
  for (int r : order) bins[SP_BINS + 1 + fill[which[r]]++] = r;          // slot order = quad order within a warp
  if (h->d_bins[side]) { cudaFree(h->d_bins[side]); h->d_bins[side] = nullptr; }
  return upload(h, &h->d_bins[side], bins.data(), bins.size());
}

int gnnuds_destroy(gnnuds_handle* h) {
  if (!h) return 0;
  cudaSetDevice(h->c.device);
  cudaStreamSynchronize(h->stream);
  void* ptrs[] = {h->d_blob, h->d_rp, h->d_ci, h->d_erp, h->d_eci, h->d_val, h->d_eval, h->d_edges, h->d_act_of_edge,
                  h->d_act_out, h->d_act_in, h->d_inc_ptr, h->d_inc_edge, h->d_inc_sign, h->d_hmin, h->d_hmax,
                  h->d_outfall, h->d_ehmax, h->d_nx, h->d_nb, h->d_ny, h->d_ne, h->d_ws, h->ro_x, h->ro_ex, h->ro_b, h->ro_a, h->d_wimg,
                  h->d_bins[0], h->d_bins[1], h->d_prp[0], h->d_prp[1], h->d_pci[0], h->d_pci[1], h->d_pval[0], h->d_pval[1], h->d_zeros, h->d_adj_act, h->d_adj_keep, h->d_arp[0], h->d_arp[1], h->d_acj[0], h->d_acj[1], h->d_aval[0], h->d_aval[1], h->d_pump, h->d_offset, h->d_pump_in, h->d_pump_out, h->d_area, h->d_edge_from, h->d_ps};
  for (void* p : ptrs) if (p) cudaFree(p);
  for (int s = 0; s < 2; ++s) for (float* p : h->d_ne_matT[s]) if (p) cudaFree(p);
  for (int s = 0; s < 2; ++s) for (uint8_t* p : h->d_ne_img[s]) if (p) cudaFree(p);
  for (int s = 0; s < 2; ++s) for (uint8_t* p : h->d_cv_img[s]) if (p) cudaFree(p);
  for (int s = 0; s < 2; ++s) for (uint8_t* p : h->d_wa[s]) if (p) cudaFree(p);
  for (int k = 0; k < 2; ++k) { if (h->pin_in[k]) cudaFreeHost(h->pin_in[k]); if (h->pin_out[k]) cudaFreeHost(h->pin_out[k]); }
  for (float* q : h->gb) if (q) cudaFree(q);
  if (h->bw_arena) cudaFree(h->bw_arena);
  for (auto& g : h->graphs) cudaGraphExecDestroy(g.second);
  for (auto& e : h->gemm_ev) { cudaEventDestroy(e.first); cudaEventDestroy(e.second); }
  if (h->ev_a) cudaEventDestroy(h->ev_a);
  if (h->ev_b) cudaEventDestroy(h->ev_b);
  for (int i = 0; i < 2; ++i) {
    if (h->ev_in[i]) cudaEventDestroy(h->ev_in[i]);
    if (h->ev_comp[i]) cudaEventDestroy(h->ev_comp[i]);
    if (h->ev_out[i]) cudaEventDestroy(h->ev_out[i]);
  }
  if (h->s_h2d) cudaStreamDestroy(h->s_h2d);
  if (h->s_side) cudaStreamDestroy(h->s_side);
  if (h->ev_fork) cudaEventDestroy(h->ev_fork);
  if (h->ev_join) cudaEventDestroy(h->ev_join);
  if (h->s_d2h) cudaStreamDestroy(h->s_d2h);
  cudaStreamDestroy(h->stream);
  delete h;
  return 0;
}

int gnnuds_set_graph(gnnuds_handle* h, const gnnuds_graph* g) {
  if (!h || !g) return GNNUDS_E_INVALID;
  const auto& c = h->c;
  CU(cudaSetDevice(c.device));
  const int n_expect = c.graph_base ? c.n_node + c.n_edge : c.n_node;
  if (g->n != n_expect) return fail(h, GNNUDS_E_INVALID, "adjacency has %d rows, expected %d", g->n, n_expect);
  if (!c.graph_base && g->m != c.n_edge) return fail(h, GNNUDS_E_INVALID, "edge adjacency has %d rows, expected %d", g->m, c.n_edge);
  if (c.conv == GNNUDS_CONV_GCN && (!g->val || (!c.graph_base && !g->e_val))) return fail(h, GNNUDS_E_INVALID, "GCN needs filter values");
  for (int i = 0; i < g->nnz; ++i) if (g->col_idx[i] < 0 || g->col_idx[i] >= g->n) return fail(h, GNNUDS_E_INVALID, "col_idx out of range");
  if (g->row_ptr[0] != 0 || g->row_ptr[g->n] != g->nnz) return fail(h, GNNUDS_E_INVALID, "row_ptr inconsistent with nnz");
  int rc;
  if ((rc = upload(h, &h->d_rp, g->row_ptr, (size_t)g->n + 1))) return rc;
  if ((rc = upload(h, &h->d_ci, g->col_idx, (size_t)g->nnz))) return rc;
  if (g->val) { if ((rc = upload(h, &h->d_val, g->val, (size_t)g->nnz))) return rc; }
  h->nnz = g->nnz; h->n_adj = g->n;
  if (!c.graph_base && (rc = upload_bins(h, 0, g->row_ptr, g->col_idx, g->val, g->n))) return rc;
  if (!c.graph_base) {
    for (int i = 0; i < g->e_nnz; ++i) if (g->e_col_idx[i] < 0 || g->e_col_idx[i] >= g->m) return fail(h, GNNUDS_E_INVALID, "e_col_idx out of range");
    if (g->e_row_ptr[0] != 0 || g->e_row_ptr[g->m] != g->e_nnz) return fail(h, GNNUDS_E_INVALID, "e_row_ptr inconsistent with e_nnz");
    if ((rc = upload(h, &h->d_erp, g->e_row_ptr, (size_t)g->m + 1))) return rc;
    if ((rc = upload(h, &h->d_eci, g->e_col_idx, (size_t)g->e_nnz))) return rc;
    if (g->e_val) { if ((rc = upload(h, &h->d_eval, g->e_val, (size_t)g->e_nnz))) return rc; }
    h->ennz = g->e_nnz;
    if ((rc = upload_bins(h, 1, g->e_row_ptr, g->e_col_idx, g->e_val, g->m))) return rc;
  }
  const int N = c.n_node, M = c.n_edge;
  h->h_edges.assign(g->edges, g->edges + 2 * (size_t)M);
  for (int i = 0; i < 2 * M; ++i) if (h->h_edges[i] < 0 || h->h_edges[i] >= N) return fail(h, GNNUDS_E_INVALID, "edge endpoint out of range");
  if ((rc = upload(h, &h->d_edges, g->edges, (size_t)2 * M))) return rc;
  std::vector<int> neg1e(M, -1), neg1n(N, -1);
  if ((rc = upload(h, &h->d_act_of_edge, g->act_of_edge ? g->act_of_edge : neg1e.data(), (size_t)M))) return rc;
  if ((rc = upload(h, &h->d_act_out, g->act_out_of_node ? g->act_out_of_node : neg1n.data(), (size_t)N))) return rc;
  if ((rc = upload(h, &h->d_act_in, g->act_in_of_node ? g->act_in_of_node : neg1n.data(), (size_t)N))) return rc;
  if (c.act && !g->act_of_edge) return fail(h, GNNUDS_E_INVALID, "act model needs act_of_edge");
  // node -> incident edges (signed incidence, base.py:432-439); self loops cancel to 0
  std::vector<int> ptr(N + 1, 0), ie, is;
  for (int m = 0; m < M; ++m) {
    const int u = h->h_edges[2 * m], v = h->h_edges[2 * m + 1];
    if (u == v) continue;
    ++ptr[u + 1]; ++ptr[v + 1];
  }
  for (int n = 0; n < N; ++n) ptr[n + 1] += ptr[n];
  ie.resize(ptr[N]); is.resize(ptr[N]);
  std::vector<int> fill(ptr.begin(), ptr.end() - 1);
  for (int m = 0; m < M; ++m) {
    const int u = h->h_edges[2 * m], v = h->h_edges[2 * m + 1];
    if (u == v) continue;
    ie[fill[u]] = m; is[fill[u]++] = +1;
    ie[fill[v]] = m; is[fill[v]++] = -1;
  }
  int dummy = 0;
  if ((rc = upload(h, &h->d_inc_ptr, ptr.data(), ptr.size()))) return rc;
  if ((rc = upload(h, &h->d_inc_edge, ie.empty() ? &dummy : ie.data(), std::max<size_t>(ie.size(), 1)))) return rc;
  if ((rc = upload(h, &h->d_inc_sign, is.empty() ? &dummy : is.data(), std::max<size_t>(is.size(), 1)))) return rc;
  if ((rc = upload(h, &h->d_hmin, g->hmin, (size_t)N))) return rc;
  if ((rc = upload(h, &h->d_hmax, g->hmax, (size_t)N))) return rc;
  if ((rc = upload(h, &h->d_outfall, g->is_outfall, (size_t)N))) return rc;
  if ((rc = upload(h, &h->d_ehmax, g->ehmax, (size_t)M))) return rc;
  // pumps, inlet offsets, pumped storage (emulator.py:590-598, 649-667)
  {
    void* old[] = {h->d_pump, h->d_offset, h->d_pump_in, h->d_pump_out, h->d_area, h->d_edge_from, h->d_ps};
    for (void* q : old) if (q) cudaFree(q);
    h->d_pump = h->d_offset = h->d_pump_in = h->d_pump_out = h->d_area = nullptr; h->d_edge_from = h->d_ps = nullptr;
    h->n_ps = 0; h->has_pump = h->has_offset = false; h->pump_all = false;
    auto any = [](const float* v, int n) { if (!v) return false; for (int i = 0; i < n; ++i) if (v[i] != 0.f) return true; return false; };
    const bool pm = any(g->pump, M) || any(g->pump_in, N) || any(g->pump_out, N), of = any(g->offset, M);
    if ((pm || of) && c.graph_base) return fail(h, GNNUDS_E_INVALID, "pumps / inlet offsets with graph_base != 0 are not supported");
    if (pm || of) {
      std::vector<int> from(M);
      for (int m = 0; m < M; ++m) from[m] = g->edges[2 * m];
      if ((rc = upload(h, &h->d_edge_from, from.data(), (size_t)M))) return rc;
    }
    if (of) { if ((rc = upload(h, &h->d_offset, g->offset, (size_t)M))) return rc; h->has_offset = true; }
    if (pm) {
      std::vector<float> zm(M, 0.f), zn(N, 0.f);
      if ((rc = upload(h, &h->d_pump, g->pump ? g->pump : zm.data(), (size_t)M))) return rc;
      if ((rc = upload(h, &h->d_pump_in, g->pump_in ? g->pump_in : zn.data(), (size_t)N))) return rc;
      if ((rc = upload(h, &h->d_pump_out, g->pump_out ? g->pump_out : zn.data(), (size_t)N))) return rc;
      if ((rc = upload(h, &h->d_area, g->area ? g->area : zn.data(), (size_t)N))) return rc;
      // ps = (area * (clip(node_edge,0,1) @ pump)) > 0: storage nodes that are the inlet of a pump
      std::vector<double> acc(N, 0.0);
      if (g->pump) for (int m = 0; m < M; ++m) if (g->edges[2 * m] != g->edges[2 * m + 1]) acc[g->edges[2 * m]] += g->pump[m];
      std::vector<int> ps;
      for (int n = 0; n < N; ++n) if (g->area && g->area[n] * acc[n] > 0) ps.push_back(n);
      h->n_ps = (int)ps.size();
      if (h->n_ps && (rc = upload(h, &h->d_ps, ps.data(), ps.size()))) return rc;
      h->has_pump = true;
      h->pump_all = g->pump != nullptr;
      for (int m = 0; m < M && h->pump_all; ++m) h->pump_all = g->pump[m] > 0.f;
    }
  }
  h->has_graph = true;
  h->has_weights = false;   // folded NodeEdge matrices depend on the incidence
  return 0;
}

int gnnuds_weight_layout(gnnuds_handle* h, int32_t* n_tensors, int64_t* sizes) {
  if (!h || !n_tensors) return GNNUDS_E_INVALID;
  *n_tensors = (int32_t)h->tens.size();
  if (sizes) for (size_t i = 0; i < h->tens.size(); ++i) sizes[i] = h->tens[i].numel;
  return 0;
}

int gnnuds_set_weights(gnnuds_handle* h, const float* blob, int64_t n_floats, const int64_t* offsets,
                       const int64_t* sizes, int32_t n_tensors) {
  if (!h || !blob || !offsets || !sizes) return GNNUDS_E_INVALID;
  if (!h->has_graph) return fail(h, GNNUDS_E_STATE, "set_graph must precede set_weights (NodeEdge folding needs the incidence)");
  const auto& c = h->c;
  CU(cudaSetDevice(c.device));
  if (n_tensors != (int)h->tens.size()) return fail(h, GNNUDS_E_INVALID, "expected %d weight tensors, got %d", (int)h->tens.size(), n_tensors);
  h->toff.resize(n_tensors);
  for (int i = 0; i < n_tensors; ++i) {
    if (sizes[i] != h->tens[i].numel) return fail(h, GNNUDS_E_INVALID, "weight tensor %d has %lld elements, expected %lld", i, (long long)sizes[i], (long long)h->tens[i].numel);
    if (offsets[i] < 0 || offsets[i] + sizes[i] > n_floats) return fail(h, GNNUDS_E_INVALID, "weight tensor %d out of blob bounds", i);
    if (offsets[i] % 4) return fail(h, GNNUDS_E_INVALID, "weight tensor %d offset not 16-byte aligned", i);
    h->toff[i] = offsets[i];
  }
  for (int64_t i = 0; i < n_floats; ++i) if (!std::isfinite(blob[i])) return fail(h, GNNUDS_E_INVALID, "non-finite weight at blob[%lld]", (long long)i);
  int rc;
  if ((rc = upload(h, &h->d_blob, blob, (size_t)n_floats))) return rc;
  if (h->tc()) {
    // bf16 K-major images (B operand = W^T) of every GEMM layer, one blob
    std::vector<int> ids;
    std::set<int> rec_ids;
    auto add = [&](const LayerW& l) { ids.push_back(l.t[0]); };
    for (int blk = 0; blk < 2; ++blk) {
      for (const SpatialW& w : h->sp[blk]) {
        if (c.graph_base) add(w.joint); else { add(w.xe); add(w.ex); add(w.conv_x); add(w.conv_e); }
      }
      for (const LayerW& l : h->tpx[blk]) { add(l); if (c.recurrent != GNNUDS_REC_CONV1D) { ids.push_back(l.t[1]); rec_ids.insert(l.t[1]); } }   // + recurrent kernel R
      for (const LayerW& l : h->tpe[blk]) { add(l); if (c.recurrent != GNNUDS_REC_CONV1D) { ids.push_back(l.t[1]); rec_ids.insert(l.t[1]); } }
    }
    add(h->resx); add(h->rese);
    for (const LayerW& l : h->flood_chain) add(l);
    h->img_off.assign(n_tensors, -1);
    std::vector<uint16_t> all;
    for (int t : ids) {
      const int Nout = h->tens[t].last();
      const int Ktot = (int)(h->tens[t].numel / Nout);
      if (Nout % 16 || Ktot % 16) return fail(h, GNNUDS_E_INVALID, "tensor-core path: weight %d is %d x %d (multiples of 16 needed)", t, Ktot, Nout);
      std::vector<uint16_t> img = pack_kmajor(blob + offsets[t], Nout, Ktot, 1, Nout, Nout, Ktot, Nout * 16);
      while (all.size() % 64) all.push_back(0);
      h->img_off[t] = (long long)all.size() * 2;
      all.insert(all.end(), img.begin(), img.end());
      if (rec_ids.count(t)) {
        // recurrent kernel R: a second image of the bf16 residual R - bf16(R) right behind the first (csrc/gru_tc.cuh)
        std::vector<float> lo((size_t)Nout * Ktot);
        for (size_t i = 0; i < lo.size(); ++i) {
          const float v = blob[offsets[t] + i];
          lo[i] = v - bf2f(f2bf(v));
        }
        std::vector<uint16_t> img2 = pack_kmajor(lo.data(), Nout, Ktot, 1, Nout, Nout, Ktot, Nout * 16);
        all.insert(all.end(), img2.begin(), img2.end());
      }
    }
    if (h->d_wimg) { cudaFree(h->d_wimg); h->d_wimg = nullptr; }
    uint16_t* d = nullptr;
    if ((rc = upload(h, &d, all.data(), all.size()))) return rc;
    h->d_wimg = reinterpret_cast<uint8_t*>(d);
  }
  if (h->nn()) {
    // zero-padded copies: flattened input widths (K) and flattened head widths (Nout) rounded up to 16
    std::vector<float> pad;
    const int N = c.n_node, M = c.n_edge, E = c.embed;
    auto r16 = [](int v) { return (v + 15) / 16 * 16; };
    auto padK = [&](const LayerW& l, int K, int Nout, gnnuds_handle::NnPad& q) {
      q.K = K; q.Kp = r16(K); q.No = q.Nop = Nout; q.w = (long long)pad.size(); q.b = -1;
      pad.insert(pad.end(), blob + offsets[l.t[0]], blob + offsets[l.t[0]] + (size_t)K * Nout);
      pad.resize(pad.size() + (size_t)(q.Kp - K) * Nout, 0.f);
    };
    auto padN = [&](const LayerW& l, int K, int Nout, gnnuds_handle::NnPad& q) {
      q.K = q.Kp = K; q.No = Nout; q.Nop = r16(Nout); q.w = (long long)pad.size();
      for (int k = 0; k < K; ++k) {
        pad.insert(pad.end(), blob + offsets[l.t[0]] + (size_t)k * Nout, blob + offsets[l.t[0]] + (size_t)(k + 1) * Nout);
        pad.resize(pad.size() + (size_t)(q.Nop - Nout), 0.f);
      }
      q.b = (long long)pad.size();
      pad.insert(pad.end(), blob + offsets[l.t[1]], blob + offsets[l.t[1]] + Nout);
      pad.resize(pad.size() + (size_t)(q.Nop - Nout), 0.f);
    };
    padK(h->emb_x, N * c.n_in, E, h->nnp_x);
    padK(h->emb_b, N * c.b_in, E / 2, h->nnp_b);
    padK(h->emb_e, M * c.e_in, E, h->nnp_e);
    if (c.act) padK(h->emb_a, M, E / 2, h->nnp_a);
    padN(h->head_out, E, c.n_out * N, h->nnp_out);
    if (c.if_flood) padN(h->head_flood, E / 2, N, h->nnp_flood);
    padN(h->head_e, E, c.e_out * M, h->nnp_eout);
    if ((rc = upload(h, &h->d_nnpad, pad.data(), pad.size()))) return rc;
  }
  // fold NodeEdge: mat = w * |inci| + b  (emulator.py:43), stored transposed + zero padded
  if (!c.graph_base && !h->nn()) {
    const int N = c.n_node, M = c.n_edge, F = c.embed / 2;
    const int IG = 256 / (F / 4);
    auto ti_for = [&](int r_out) { int t = (r_out + IG - 1) / IG; t = (t + 3) / 4 * 4; return t; };
    h->ti_n = ti_for(N); h->ti_m = ti_for(M);
    h->ldm_n = IG * h->ti_n; h->ldm_m = IG * h->ti_m;
    for (int s = 0; s < 2; ++s) { for (float* p : h->d_ne_matT[s]) if (p) cudaFree(p); h->d_ne_matT[s].clear(); }
    std::vector<float> inci((size_t)N * M, 0.f);
    for (int m = 0; m < M; ++m) {
      const int u = h->h_edges[2 * m], v = h->h_edges[2 * m + 1];
      inci[(size_t)u * M + m] += 1.f; inci[(size_t)v * M + m] -= 1.f;
    }
    for (auto& v : inci) v = std::fabs(v);
    for (int blk = 0; blk < 2; ++blk)
      for (int l = 0; l < c.n_sp_layer; ++l) {
        const SpatialW& w = h->sp[blk][l];
        {  // N side: mat (N x M); matT[j=m][i=n]
          const float* wt = blob + offsets[w.ne_n.t[0]]; const float* bs = blob + offsets[w.ne_n.t[1]];
          std::vector<float> mt((size_t)M * h->ldm_n, 0.f);
          for (int n = 0; n < N; ++n) for (int m = 0; m < M; ++m)
            mt[(size_t)m * h->ldm_n + n] = wt[(size_t)n * M + m] * inci[(size_t)n * M + m] + bs[(size_t)n * M + m];
          float* d = nullptr;
          if ((rc = upload(h, &d, mt.data(), mt.size()))) return rc;
          h->d_ne_matT[0].push_back(d);
        }
        {  // M side: mat (M x N) with |inci|^T; matT[j=n][i=m]
          const float* wt = blob + offsets[w.ne_m.t[0]]; const float* bs = blob + offsets[w.ne_m.t[1]];
          std::vector<float> mt((size_t)N * h->ldm_m, 0.f);
          for (int m = 0; m < M; ++m) for (int n = 0; n < N; ++n)
            mt[(size_t)n * h->ldm_m + m] = wt[(size_t)m * N + n] * inci[(size_t)n * M + m] + bs[(size_t)m * N + n];
          float* d = nullptr;
          if ((rc = upload(h, &d, mt.data(), mt.size()))) return rc;
          h->d_ne_matT[1].push_back(d);
        }
      }
    if (h->tc()) {
      // the same folded matrices as bf16 K-major images (A operand of the NodeEdge MMA in sp_conv_kernel)
      for (int s = 0; s < 2; ++s) { for (uint8_t* p : h->d_ne_img[s]) if (p) cudaFree(p); h->d_ne_img[s].clear(); }
      const int Rs[2] = {N, M}, Ks[2] = {M, N};
      const int g = h->sp_g;
      for (int s = 0; s < 2; ++s) {
        h->ne_kp[s] = (g * Ks[s] + 15) / 16 * 16;
        h->ne_pstride[s] = (uint32_t)((g * Rs[s] + 7) / 8 * 8) * 16u;
        h->ne_img_bytes[s] = (uint32_t)(h->ne_kp[s] / 8) * h->ne_pstride[s];
      }
      for (int blk = 0; blk < 2; ++blk)
        for (int l = 0; l < c.n_sp_layer; ++l) {
          const SpatialW& w = h->sp[blk][l];
          for (int s = 0; s < 2; ++s) {
            const LayerW& lw = s == 0 ? w.ne_n : w.ne_m;
            const float* wt = blob + offsets[lw.t[0]]; const float* bs = blob + offsets[lw.t[1]];
            const int Rg = g * Rs[s], Kg = g * Ks[s];
            std::vector<float> md((size_t)Rg * Kg, 0.f);       // block diagonal: one copy per sample of the group
            for (int q = 0; q < g; ++q)
              for (int r = 0; r < Rs[s]; ++r)
                for (int k = 0; k < Ks[s]; ++k) {
                  const float inc = s == 0 ? inci[(size_t)r * M + k] : inci[(size_t)k * M + r];
                  md[(size_t)(q * Rs[s] + r) * Kg + q * Ks[s] + k] = wt[(size_t)r * Ks[s] + k] * inc + bs[(size_t)r * Ks[s] + k];
                }
            std::vector<uint16_t> img = pack_kmajor(md.data(), Rg, Kg, Kg, 1, 0, h->ne_kp[s], (int)h->ne_pstride[s]);
            uint16_t* d = nullptr;
            if ((rc = upload(h, &d, img.data(), img.size()))) return rc;
            h->d_ne_img[s].push_back(reinterpret_cast<uint8_t*>(d));
          }
        }
      // GAT: conv weight images with 16 extra output columns [hi(W a_self) lo(W a_self) hi(W a_neigh) lo(W a_neigh) 0..]
      // (hi/lo = 16-bit split of the fp32 product), so sp_agg_kernel reads the attention scores as accumulator columns
      for (int s = 0; s < 2; ++s) { for (uint8_t* p : h->d_cv_img[s]) if (p) cudaFree(p); h->d_cv_img[s].clear(); h->cv_img_bytes[s].clear(); }
      for (int s = 0; s < 2; ++s) { for (uint8_t* p : h->d_wa[s]) if (p) cudaFree(p); h->d_wa[s].clear(); }
      if (c.conv == GNNUDS_CONV_GAT && c.embed % 16 == 0)
        for (int blk = 0; blk < 2; ++blk)
          for (int l = 0; l < c.n_sp_layer; ++l) {
            const SpatialW& w = h->sp[blk][l];
            for (int s = 0; s < 2; ++s) {
              const LayerW& lw = s == 0 ? w.conv_x : w.conv_e;
              const int E = c.embed, NW = E + 16, Ktot = (int)(h->tens[lw.t[0]].numel / E);
              const float* Wk = blob + offsets[lw.t[0]]; const float* as = blob + offsets[lw.t[1]]; const float* an = blob + offsets[lw.t[2]];
              // sp_aggt_kernel: the same four vectors as 8-row A operand panels [K/8][8 rows][8 K], rows 0..3
              std::vector<float> wx((size_t)Ktot * NW, 0.f);
              std::vector<uint16_t> wav((size_t)Ktot * 8, 0);
              for (int k = 0; k < Ktot; ++k) {
                double ds = 0.0, dn = 0.0;
                for (int n = 0; n < E; ++n) { wx[(size_t)k * NW + n] = Wk[(size_t)k * E + n]; ds += (double)Wk[(size_t)k * E + n] * as[n]; dn += (double)Wk[(size_t)k * E + n] * an[n]; }
                const float fs = (float)ds, fn = (float)dn, fsh = bf2f(f2bf(fs)), fnh = bf2f(f2bf(fn));
                wx[(size_t)k * NW + E] = fsh; wx[(size_t)k * NW + E + 1] = fs - fsh;
                wx[(size_t)k * NW + E + 2] = fnh; wx[(size_t)k * NW + E + 3] = fn - fnh;
                const float four[4] = {fsh, fs - fsh, fnh, fn - fnh};
                for (int r = 0; r < 4; ++r) wav[(size_t)(k >> 3) * 64 + (size_t)r * 8 + (k & 7)] = f2bf(four[r]);
              }
              { uint16_t* dw = nullptr; if ((rc = upload(h, &dw, wav.data(), wav.size()))) return rc; h->d_wa[s].push_back(reinterpret_cast<uint8_t*>(dw)); }
              std::vector<uint16_t> img = pack_kmajor(wx.data(), NW, Ktot, 1, NW, NW, Ktot, NW * 16);
              uint16_t* d = nullptr;
              if ((rc = upload(h, &d, img.data(), img.size()))) return rc;
              h->d_cv_img[s].push_back(reinterpret_cast<uint8_t*>(d));
              h->cv_img_bytes[s].push_back((uint32_t)(img.size() * 2));
            }
          }
    }
  }
  h->has_weights = true;
  return 0;
}

int gnnuds_set_norm(gnnuds_handle* h, const float* norm_x, int32_t c_x, const float* norm_b, int32_t c_b,
                    const float* norm_y, int32_t c_y, const float* norm_e, int32_t c_e) {
  if (!h || !norm_x || !norm_b || !norm_y || !norm_e) return GNNUDS_E_INVALID;
  const auto& c = h->c;
  CU(cudaSetDevice(c.device));
  if (c_x < c.n_in || c_b < c.b_in || c_e < c.e_in) return fail(h, GNNUDS_E_INVALID, "norm arrays have fewer channels than the inputs");
  if (c_y < 3 + (c.if_flood ? 1 : 0)) return fail(h, GNNUDS_E_INVALID, "norm_y needs >= %d channels", 3 + (c.if_flood ? 1 : 0));
  int rc;
  if ((rc = upload(h, &h->d_nx, norm_x, (size_t)2 * c.n_node * c_x))) return rc;
  if ((rc = upload(h, &h->d_nb, norm_b, (size_t)2 * c.n_node * c_b))) return rc;
  if ((rc = upload(h, &h->d_ny, norm_y, (size_t)2 * c.n_node * c_y))) return rc;
  if ((rc = upload(h, &h->d_ne, norm_e, (size_t)2 * c.n_edge * c_e))) return rc;
  h->cx_n = c_x; h->cb_n = c_b; h->cy_n = c_y; h->ce_n = c_e;
  h->has_norm = true;
  return 0;
}

int gnnuds_set_post_proc_variant(gnnuds_handle* h, int32_t variant) {
  if (!h) return GNNUDS_E_INVALID;
  if (variant != 0 && variant != 1) return fail(h, GNNUDS_E_INVALID, "post-processing variant %d unknown (0 numpy, 1 tf)", variant);
  h->pp_variant = variant;
  return 0;
}

int gnnuds_set_adj_action(gnnuds_handle* h, const int32_t* act_entry, const uint8_t* base_keep) {
  if (!h || !act_entry || !base_keep) return GNNUDS_E_INVALID;
  const auto& c = h->c;
  if (!h->has_graph) return fail(h, GNNUDS_E_STATE, "set_adj_action: set the graph first");
  if (!c.act) return fail(h, GNNUDS_E_INVALID, "use_adj needs an act model (emulator.py:178)");
  if (c.conv != GNNUDS_CONV_GAT)
    return fail(h, GNNUDS_E_INVALID, "use_adj is supported for GAT only: with GCN the reference hands a 4-D adjacency to spektral's "
                                     "gcn_filter, which raises (np.diag_indices_from needs equal dimensions)");
  CU(cudaSetDevice(c.device));
  int rc;
  std::vector<int> ae(act_entry, act_entry + h->nnz);
  for (int v : ae) if (v < 0 || v > c.n_act) return fail(h, GNNUDS_E_INVALID, "set_adj_action: controlled-link index %d out of range", v);
  if ((rc = upload(h, &h->d_adj_act, ae.data(), ae.size()))) return rc;
  if ((rc = upload(h, &h->d_adj_keep, base_keep, (size_t)h->nnz))) return rc;
  h->use_adj = true;
  return 0;
}

int gnnuds_tc_operand_is_half(void) { return GU_F16 ? 16 : 0; }

int gnnuds_measure_simt_tflops(gnnuds_handle* h, float* tflops) {
  if (!h || !tflops) return GNNUDS_E_INVALID;
  CU(cudaSetDevice(h->c.device));
  const int blocks = h->sm_count * 8, threads = 256, iters = 1 << 14;
  float* d = nullptr;
  CU(cudaMalloc(&d, (size_t)blocks * threads * sizeof(float)));
  cudaEvent_t e0, e1;
  CU(cudaEventCreate(&e0)); CU(cudaEventCreate(&e1));
  float best = 0.f;
  for (int rep = 0; rep < 4; ++rep) {
    CU(cudaEventRecord(e0, h->stream));
    fp32_peak_kernel<<<blocks, threads, 0, h->stream>>>(d, iters);
    CU(cudaEventRecord(e1, h->stream));
    CU(cudaEventSynchronize(e1));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, e0, e1));
    if (rep && ms > 0.f) best = std::max(best, (float)(2.0 * 16.0 * iters * blocks * threads / (ms * 1e-3) / 1e12));
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
  *tflops = best;
  return 0;
}

int gnnuds_output_dims(const gnnuds_handle* h, int32_t* y_channels, int32_t* ey_channels) {
  if (!h) return GNNUDS_E_INVALID;
  if (y_channels) *y_channels = h->ny();
  if (ey_channels) *ey_channels = h->c.e_out;
  return 0;
}

int gnnuds_predict(gnnuds_handle* h, const float* x, const float* b, const float* a, const float* ex, int64_t B,
                   float* y, float* ey, int32_t where) {
  int rc = check_ready(h);
  if (rc) return rc;
  const auto& c = h->c;
  if (!x || !b || !ex || !y || !ey || B < 0) return fail(h, GNNUDS_E_INVALID, "null buffer");
  if (c.act && !a) return fail(h, GNNUDS_E_INVALID, "act model needs settings `a`");
  CU(cudaSetDevice(c.device));
  if (B == 0) return 0;
  int64_t Bc = pick_chunk(h, B);
  if (where == GNNUDS_MEM_HOST && c.max_chunk <= 0 && h->nn()) {
    // host buffers, NN-cat: the copies take longer than the compute and only overlap it across chunks, so prefer ~8
    // chunks as long as one chunk still fills the machine (>= 4 waves of 128-row tiles).  The conv models are
    // compute-dominated and lose more to short launches than the overlap returns (C3b: 118 -> 101 M/s when split).
    const int64_t rows_per_sc = (int64_t)std::max(c.seq_in, c.seq_out);
    const int64_t fill = ((int64_t)h->sm_count * 128 * 4 + rows_per_sc - 1) / rows_per_sc;
    Bc = std::min(Bc, std::max((B + 7) / 8, fill));
  }
  if ((rc = ensure_workspace(h, Bc))) return rc;
  const int N = c.n_node, M = c.n_edge, Ti = c.seq_in, To = c.seq_out;
  const int64_t sx = (int64_t)Ti * N * c.n_in, sb = (int64_t)To * N * c.b_in, sa = (int64_t)To * c.n_act,
                se = (int64_t)Ti * M * c.e_in, sy = (int64_t)To * N * h->ny(), sey = (int64_t)To * M * c.e_out;
  begin_timing(h);
  if (h->profiling) CU(cudaEventRecord(h->ev_a, h->stream));
  const bool staged = where == GNNUDS_MEM_HOST && !h->debug_stop && !(getenv("GNNUDS_NO_PIN_RING") && getenv("GNNUDS_NO_PIN_RING")[0] == '1') &&
                      (is_pageable(x) || is_pageable(b) || is_pageable(ex) || (a && is_pageable(a)) || is_pageable(y) || is_pageable(ey));
  if (staged) {
    // Pageable caller buffers (plain NumPy arrays): cudaMemcpyAsync from pageable memory blocks the host and runs at a
    // fraction of the pinned rate, so chunks go through a ring of two pinned input / output sets.  Worker threads copy
    // caller memory <-> ring while the GPU computes: stage-in of chunk i+1 and drain of chunk i-1 overlap chunk i.
    if ((rc = ensure_pinned(h, (size_t)(Bc * (sx + sb + sa + se)), (size_t)(Bc * (sy + sey))))) return rc;
    float* sin[2][4] = {{h->in_x, h->in_b, h->in_a, h->in_ex}, {h->in2_x, h->in2_b, h->in2_a, h->in2_ex}};
    float* sout[2][2] = {{h->out_y, h->out_ey}, {h->out2_y, h->out2_ey}};
    CU(cudaEventRecord(h->ev_comp[0], h->stream));
    CU(cudaStreamWaitEvent(h->s_h2d, h->ev_comp[0], 0));
    int64_t i = 0, prev_b0 = 0, prev_nb = 0;
    auto drain = [&](int k, int64_t b0, int64_t nb) -> int {
      CU(cudaEventSynchronize(h->ev_out[k]));
      parallel_copy({{y + b0 * sy, h->pin_out[k], (size_t)(nb * sy) * sizeof(float)},
                     {ey + b0 * sey, h->pin_out[k] + nb * sy, (size_t)(nb * sey) * sizeof(float)}});
      return 0;
    };
    for (int64_t b0 = 0; b0 < B; b0 += Bc, ++i) {
      const int64_t nb = std::min(Bc, B - b0);
      const int k = (int)(i & 1);
      if (i >= 2) CU(cudaEventSynchronize(h->ev_in[k]));                    // the H2D copy of chunk i-2 has left this ring slot
      float *qx = h->pin_in[k], *qb = qx + nb * sx, *qe = qb + nb * sb, *qa = qe + nb * se;
      parallel_copy({{qx, x + b0 * sx, (size_t)(nb * sx) * sizeof(float)}, {qb, b + b0 * sb, (size_t)(nb * sb) * sizeof(float)},
                     {qe, ex + b0 * se, (size_t)(nb * se) * sizeof(float)}, {qa, a ? a + b0 * sa : nullptr, a ? (size_t)(nb * sa) * sizeof(float) : 0}});
      if (i >= 2) CU(cudaStreamWaitEvent(h->s_h2d, h->ev_comp[k], 0));
      CU(cudaMemcpyAsync(sin[k][0], qx, nb * sx * sizeof(float), cudaMemcpyHostToDevice, h->s_h2d));
      CU(cudaMemcpyAsync(sin[k][1], qb, nb * sb * sizeof(float), cudaMemcpyHostToDevice, h->s_h2d));
      CU(cudaMemcpyAsync(sin[k][3], qe, nb * se * sizeof(float), cudaMemcpyHostToDevice, h->s_h2d));
      if (a) CU(cudaMemcpyAsync(sin[k][2], qa, nb * sa * sizeof(float), cudaMemcpyHostToDevice, h->s_h2d));
      CU(cudaEventRecord(h->ev_in[k], h->s_h2d));
      CU(cudaStreamWaitEvent(h->stream, h->ev_in[k], 0));
      if (i >= 2) CU(cudaStreamWaitEvent(h->stream, h->ev_out[k], 0));
      if ((rc = forward_any(h, sin[k][0], sin[k][1], a ? sin[k][2] : nullptr, sin[k][3], nb, sout[k][0], sout[k][1], false))) return rc;
      CU(cudaEventRecord(h->ev_comp[k], h->stream));
      CU(cudaStreamWaitEvent(h->s_d2h, h->ev_comp[k], 0));
      CU(cudaMemcpyAsync(h->pin_out[k], sout[k][0], nb * sy * sizeof(float), cudaMemcpyDeviceToHost, h->s_d2h));
      CU(cudaMemcpyAsync(h->pin_out[k] + nb * sy, sout[k][1], nb * sey * sizeof(float), cudaMemcpyDeviceToHost, h->s_d2h));
      CU(cudaEventRecord(h->ev_out[k], h->s_d2h));
      if (i >= 1 && (rc = drain(k ^ 1, prev_b0, prev_nb))) return rc;
      prev_b0 = b0; prev_nb = nb;
    }
    if (i >= 1 && (rc = drain((int)((i - 1) & 1), prev_b0, prev_nb))) return rc;
  } else if (where == GNNUDS_MEM_HOST && !h->debug_stop) {
    // Host buffers: three streams and two staging sets.  The H2D copy of chunk i+1 and the D2H copy of
    // chunk i-1 run while chunk i computes; events order the reuse of a staging set (i and i+2 share one).
    float* sin[2][4] = {{h->in_x, h->in_b, h->in_a, h->in_ex}, {h->in2_x, h->in2_b, h->in2_a, h->in2_ex}};
    float* sout[2][2] = {{h->out_y, h->out_ey}, {h->out2_y, h->out2_ey}};
    CU(cudaEventRecord(h->ev_comp[0], h->stream));     // earlier work on the handle's stream (workspace memset)
    CU(cudaStreamWaitEvent(h->s_h2d, h->ev_comp[0], 0));
    int64_t i = 0;
    for (int64_t b0 = 0; b0 < B; b0 += Bc, ++i) {
      const int64_t nb = std::min(Bc, B - b0);
      const int k = (int)(i & 1);
      const float *px = x + b0 * sx, *pb = b + b0 * sb, *pa = a ? a + b0 * sa : nullptr, *pe = ex + b0 * se;
      if (i >= 2) CU(cudaStreamWaitEvent(h->s_h2d, h->ev_comp[k], 0));       // chunk i-2 has consumed this input set
      CU(cudaMemcpyAsync(sin[k][0], px, nb * sx * sizeof(float), cudaMemcpyHostToDevice, h->s_h2d));
      CU(cudaMemcpyAsync(sin[k][1], pb, nb * sb * sizeof(float), cudaMemcpyHostToDevice, h->s_h2d));
      CU(cudaMemcpyAsync(sin[k][3], pe, nb * se * sizeof(float), cudaMemcpyHostToDevice, h->s_h2d));
      if (pa) CU(cudaMemcpyAsync(sin[k][2], pa, nb * sa * sizeof(float), cudaMemcpyHostToDevice, h->s_h2d));
      CU(cudaEventRecord(h->ev_in[k], h->s_h2d));
      CU(cudaStreamWaitEvent(h->stream, h->ev_in[k], 0));
      if (i >= 2) CU(cudaStreamWaitEvent(h->stream, h->ev_out[k], 0));       // chunk i-2's results have left this output set
      if ((rc = forward_any(h, sin[k][0], sin[k][1], pa ? sin[k][2] : nullptr, sin[k][3], nb, sout[k][0], sout[k][1], false))) return rc;
      CU(cudaEventRecord(h->ev_comp[k], h->stream));
      CU(cudaStreamWaitEvent(h->s_d2h, h->ev_comp[k], 0));
      CU(cudaMemcpyAsync(y + b0 * sy, sout[k][0], nb * sy * sizeof(float), cudaMemcpyDeviceToHost, h->s_d2h));
      CU(cudaMemcpyAsync(ey + b0 * sey, sout[k][1], nb * sey * sizeof(float), cudaMemcpyDeviceToHost, h->s_d2h));
      CU(cudaEventRecord(h->ev_out[k], h->s_d2h));
    }
    for (int k = 0; k < 2 && k < i; ++k) CU(cudaStreamWaitEvent(h->stream, h->ev_out[k], 0));
  } else {
    for (int64_t b0 = 0; b0 < B; b0 += Bc) {
      const int64_t nb = std::min(Bc, B - b0);
      const float *px = x + b0 * sx, *pb = b + b0 * sb, *pa = a ? a + b0 * sa : nullptr, *pe = ex + b0 * se;
      float *py = y + b0 * sy, *pey = ey + b0 * sey;
      if (where == GNNUDS_MEM_HOST) {     // debug_stop: stage serially, nothing to copy back
        CU(cudaMemcpyAsync(h->in_x, px, nb * sx * sizeof(float), cudaMemcpyHostToDevice, h->stream));
        CU(cudaMemcpyAsync(h->in_b, pb, nb * sb * sizeof(float), cudaMemcpyHostToDevice, h->stream));
        CU(cudaMemcpyAsync(h->in_ex, pe, nb * se * sizeof(float), cudaMemcpyHostToDevice, h->stream));
        if (pa) CU(cudaMemcpyAsync(h->in_a, pa, nb * sa * sizeof(float), cudaMemcpyHostToDevice, h->stream));
        if ((rc = forward_any(h, h->in_x, h->in_b, pa ? h->in_a : nullptr, h->in_ex, nb, h->out_y, h->out_ey, false))) return rc;
      } else {
        if ((rc = forward_any(h, px, pb, pa, pe, nb, py, pey, false))) return rc;
      }
      if (h->debug_stop) break;
    }
  }
  if (h->profiling) CU(cudaEventRecord(h->ev_b, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  if (h->profiling) CU(cudaEventElapsedTime(&h->last_total_ms, h->ev_a, h->ev_b));
  return finish_timing(h);
}

int gnnuds_simulate(gnnuds_handle* h, const float* states, const float* runoff, int32_t seq_r, const float* a,
                    const float* edge_states, int64_t n_windows, float* y, float* ey, int32_t where) {
  int rc = check_ready(h);
  if (rc) return rc;
  const auto& c = h->c;
  if (seq_r < c.seq_out) return fail(h, GNNUDS_E_INVALID, "runoff windows shorter than seq_out");
  if (seq_r == c.seq_out) return gnnuds_predict(h, states, runoff, a, edge_states, n_windows, y, ey, where);
  // runoff (T', seq_r, N, b_in): keep the first seq_out steps of every window (emulator.py:532)
  if (where != GNNUDS_MEM_HOST) return fail(h, GNNUDS_E_INVALID, "simulate with seq_r > seq_out takes host buffers");
  const size_t row = (size_t)c.n_node * c.b_in;
  std::vector<float> r((size_t)n_windows * c.seq_out * row);
  for (int64_t w = 0; w < n_windows; ++w)
    std::memcpy(r.data() + (size_t)w * c.seq_out * row, runoff + (size_t)w * seq_r * row, (size_t)c.seq_out * row * sizeof(float));
  return gnnuds_predict(h, states, r.data(), a, edge_states, n_windows, y, ey, where);
}

int gnnuds_rollout(gnnuds_handle* h, int32_t rule, const float* x0, const float* b, const float* a, const float* ex0,
                   int64_t B, int32_t n_chunks, float* y, float* ey, int32_t where) {
  int rc = check_ready(h);
  if (rc) return rc;
  const auto& c = h->c;
  if (rule < GNNUDS_ROLL_MPC || rule > GNNUDS_ROLL_NORM) return fail(h, GNNUDS_E_INVALID, "unknown feedback rule %d", rule);
  if (c.seq_in != c.seq_out) return fail(h, GNNUDS_E_INVALID, "rollout needs seq_in == seq_out (mpc.py:573-579 feeds back seq_out steps)");
  if (rule == GNNUDS_ROLL_MPC && !c.if_flood) return fail(h, GNNUDS_E_INVALID, "MPC feedback rule needs if_flood (mpc.py:575 drops two channels)");
  if (!x0 || !b || !ex0 || !y || !ey || n_chunks < 1) return fail(h, GNNUDS_E_INVALID, "bad argument");
  if (c.act && !a) return fail(h, GNNUDS_E_INVALID, "act model needs settings `a`");
  CU(cudaSetDevice(c.device));
  if (B == 0) return 0;
  const bool norm_io = rule == GNNUDS_ROLL_NORM;
  const int N = c.n_node, M = c.n_edge, T = c.seq_out;
  const int nyc = norm_io ? 3 + (c.if_flood ? 1 : 0) : h->ny();
  const int64_t Bc = pick_chunk(h, B);
  if ((rc = ensure_workspace(h, Bc))) return rc;
  // per-chunk device state: x, ex (one window), full-horizon b / a / y / ey slices
  const int64_t Th = (int64_t)n_chunks * T;
  const int64_t sx = (int64_t)T * N * c.n_in, se = (int64_t)T * M * c.e_in;
  const int64_t sbh = Th * N * c.b_in, sah = Th * c.n_act, syh = Th * N * nyc, seyh = Th * M * c.e_out;
  // horizon staging (host callers) and per-chunk state: grow-only buffers of the handle
  float *d_b = nullptr, *d_a = nullptr, *d_y = nullptr, *d_ey = nullptr;
  if (where == GNNUDS_MEM_HOST) {
    if ((rc = grow(h, 0, (size_t)(Bc * sbh), &d_b)) || (rc = grow(h, 2, (size_t)(Bc * syh), &d_y)) || (rc = grow(h, 3, (size_t)(Bc * seyh), &d_ey))) return rc;
    if (a && (rc = grow(h, 1, (size_t)(Bc * sah), &d_a))) return rc;
  }
  if ((rc = ensure_rollout_state(h, Bc))) return rc;
  begin_timing(h);
  if (h->profiling) cudaEventRecord(h->ev_a, h->stream);
  rc = 0;
  // every autoregressive loop of the reference runs the TF variant of the post-processing (mpc.py:575 and mbrl.py:337
  // call predict_tf, emulator.py:413 calls post_proc_tf)
  const int saved_variant = h->pp_variant;
  h->pp_variant = 1;
  struct Restore { gnnuds_handle* h; int v; ~Restore() { h->pp_variant = v; } } restore{h, saved_variant};
  for (int64_t b0 = 0; b0 < B && !rc; b0 += Bc) {
    const int64_t nb = std::min(Bc, B - b0);
    const float *hb = b + b0 * sbh, *ha = a ? a + b0 * sah : nullptr;
    float *hy = y + b0 * syh, *hey = ey + b0 * seyh;
    const float *db, *da;
    float *dy, *dey;
    cudaMemcpyKind kin = where == GNNUDS_MEM_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice;
    cudaMemcpyAsync(h->ro_x, x0 + b0 * sx, nb * sx * sizeof(float), kin, h->stream);
    cudaMemcpyAsync(h->ro_ex, ex0 + b0 * se, nb * se * sizeof(float), kin, h->stream);
    if (where == GNNUDS_MEM_HOST) {
      cudaMemcpyAsync(d_b, hb, nb * sbh * sizeof(float), kin, h->stream);
      if (ha) cudaMemcpyAsync(d_a, ha, nb * sah * sizeof(float), kin, h->stream);
      db = d_b; da = ha ? d_a : nullptr; dy = d_y; dey = d_ey;
    } else {
      db = hb; da = ha; dy = hy; dey = hey;
    }
    // the chunk loop - slices, n_chunks forwards, scatters, feedback - touches device buffers only: one cached CUDA graph
    // per shape (a roll of ONE scenario is latency-bound: ~45 kernels per chunk)
    if ((rc = side_prepare(h, nb))) break;
    char key[256];
    snprintf(key, sizeof key, "roll|%lld|%d|%d|%d|%p|%p|%p|%p", (long long)nb, n_chunks, rule, h->side_on ? 1 : 0, (const void*)db, (const void*)da,
             (void*)dy, (void*)dey);
    auto chunk_loop = [&]() -> int {
      int rc = 0;
    for (int k = 0; k < n_chunks && !rc; ++k) {
        // slice chunk k of b / a into compact (nb,T,...) buffers
        const long long Lb = (long long)N * c.b_in;
        slice_time_kernel<<<grid_for(h, nb * T * Lb, 256, 16), 256, 0, h->stream>>>(db, h->ro_b, nb, (int)Th, k * T, T, Lb, 0);
        ++h->launches;
        if (da) {
          slice_time_kernel<<<grid_for(h, nb * T * c.n_act, 256, 16), 256, 0, h->stream>>>(da, h->ro_a, nb, (int)Th, k * T, T, c.n_act, 0);
          ++h->launches;
        }
        rc = forward_any(h, h->ro_x, h->ro_b, da ? h->ro_a : nullptr, h->ro_ex, nb, h->out_y, h->out_ey, norm_io);
        if (rc) break;
        // scatter outputs into the horizon tensors
        slice_time_kernel<<<grid_for(h, nb * T * (long long)N * nyc, 256, 16), 256, 0, h->stream>>>(h->out_y, dy, nb, (int)Th, k * T, T, (long long)N * nyc, 1);
        slice_time_kernel<<<grid_for(h, nb * T * (long long)M * c.e_out, 256, 16), 256, 0, h->stream>>>(h->out_ey, dey, nb, (int)Th, k * T, T, (long long)M * c.e_out, 1);
        h->launches += 2;
        if (k + 1 < n_chunks) {
          FeedP p{};
          p.y = h->out_y; p.ny = nyc; p.ey = h->out_ey; p.e_out = c.e_out; p.b = h->ro_b; p.b_in = c.b_in;
          p.a = da ? h->ro_a : nullptr; p.n_act = c.n_act; p.act_of_edge = h->d_act_of_edge;
          p.x = h->ro_x; p.n_in = c.n_in; p.ex = h->ro_ex; p.e_in = c.e_in;
          p.S = nb * T; p.N = N; p.M = M; p.if_flood = c.if_flood ? 1 : 0; p.rule = rule;
          feedback_kernel<<<grid_for(h, p.S * (N + M), 256, 16), 256, 0, h->stream>>>(p);
          ++h->launches;
        }
        if (cudaGetLastError() != cudaSuccess) rc = fail(h, GNNUDS_E_CUDA, "rollout kernel launch failed");
      }
      return rc;
    };
    rc = run_graphed(h, key, chunk_loop);
    if (!rc && where == GNNUDS_MEM_HOST) {
      cudaMemcpyAsync(hy, d_y, nb * syh * sizeof(float), cudaMemcpyDeviceToHost, h->stream);
      cudaMemcpyAsync(hey, d_ey, nb * seyh * sizeof(float), cudaMemcpyDeviceToHost, h->stream);
    }
  }
  if (h->profiling) cudaEventRecord(h->ev_b, h->stream);
  cudaError_t e = cudaStreamSynchronize(h->stream);
  if (rc) return rc;
  if (e != cudaSuccess) return fail(h, GNNUDS_E_CUDA, "rollout: %s", cudaGetErrorString(e));
  if (h->profiling) cudaEventElapsedTime(&h->last_total_ms, h->ev_a, h->ev_b);
  return finish_timing(h);
}

int gnnuds_objective(gnnuds_handle* h, const float* y, const float* q_in0, int64_t B, int32_t T, const float* flood_w,
                     const float* wwtp_w, const float* rough_w, const float* gamma, float* obj, int32_t where) {
  if (!h || !y || !q_in0 || !flood_w || !wwtp_w || !obj) return GNNUDS_E_INVALID;
  const auto& c = h->c;
  CU(cudaSetDevice(c.device));
  if (B == 0) return 0;
  const int N = c.n_node, nyc = h->ny();
  float *dfw = nullptr, *dww = nullptr, *drw = nullptr, *dg = nullptr, *dy = nullptr, *dq = nullptr, *dobj = nullptr;
  int rc = 0;
  auto up = [&](int slot, float** d, const float* src, size_t n) {
    int r = grow(h, slot, n, d);
    if (r) return r;
    return cudaMemcpyAsync(*d, src, n * sizeof(float), cudaMemcpyHostToDevice, h->stream) != cudaSuccess ? fail(h, GNNUDS_E_CUDA, "objective: copy failed") : 0;
  };
  if ((rc = up(4, &dfw, flood_w, N)) || (rc = up(5, &dww, wwtp_w, N))) return rc;
  if (rough_w && (rc = up(6, &drw, rough_w, N))) return rc;
  if (gamma && (rc = up(7, &dg, gamma, T))) return rc;
  const float *py = y, *pq = q_in0;
  float* po = obj;
  if (where == GNNUDS_MEM_HOST) {
    if ((rc = up(8, &dy, y, (size_t)B * T * N * nyc)) || (rc = up(9, &dq, q_in0, (size_t)B * N)) || (rc = grow(h, 10, (size_t)B, &dobj))) return rc;
    py = dy; pq = dq; po = dobj;
  }
  ObjP p{py, nyc, pq, dfw, dww, drw, dg, po, B, T, N};
  objective_kernel<<<grid_for(h, B, 1, 16), 128, 0, h->stream>>>(p);
  ++h->launches;
  if (where == GNNUDS_MEM_HOST) cudaMemcpyAsync(obj, dobj, B * sizeof(float), cudaMemcpyDeviceToHost, h->stream);
  cudaError_t e = cudaStreamSynchronize(h->stream);
  if (e != cudaSuccess) return fail(h, GNNUDS_E_CUDA, "objective: %s", cudaGetErrorString(e));
  return 0;
}

int gnnuds_evaluate(gnnuds_handle* h, int32_t rule, const float* x0, const float* ex0, const float* b, int32_t n_b,
                    const float* dec, int32_t n_step, int32_t r_step, int64_t B, int32_t n_chunks, const float* flood_w,
                    const float* wwtp_w, const float* rough_w, const float* gamma, float* obj, int32_t where, int32_t obj_where) {
  int rc = check_ready(h);
  if (rc) return rc;
  const auto& c = h->c;
  if (rule < GNNUDS_ROLL_MPC || rule > GNNUDS_ROLL_MBRL) return fail(h, GNNUDS_E_INVALID, "evaluate: feedback rule %d (MPC or MBRL)", rule);
  if (!x0 || !ex0 || !b || !dec || !flood_w || !wwtp_w || !obj || n_chunks < 1 || n_b < 1 || n_step < 1 || r_step < 1 || B < 0)
    return fail(h, GNNUDS_E_INVALID, "evaluate: bad argument");
  if (!c.act || !c.if_flood) return fail(h, GNNUDS_E_INVALID, "evaluate needs an act + if_flood model (mpc.py:575, astlingen.py:75-99)");
  if (n_chunks > 1 && c.seq_in != c.seq_out) return fail(h, GNNUDS_E_INVALID, "chunked evaluation needs seq_in == seq_out (mpc.py:573-579)");
  CU(cudaSetDevice(c.device));
  if (B == 0) return 0;
  const int N = c.n_node, M = c.n_edge, T = c.seq_out, Ti = c.seq_in, nyc = h->ny();
  const int64_t Th = (int64_t)n_chunks * T;
  const int64_t Bc = pick_chunk(h, B);
  if ((rc = ensure_workspace(h, Bc)) || (rc = ensure_rollout_state(h, Bc))) return rc;
  const size_t nx0 = (size_t)Ti * N * c.n_in, nex0 = (size_t)Ti * M * c.e_in, nbb = (size_t)n_b * Th * N * c.b_in,
               ndec = (size_t)B * n_step * c.n_act;
  const float *dx0 = x0, *dex0 = ex0, *db = b, *ddec = dec, *dfw = flood_w, *dww = wwtp_w, *drw = rough_w, *dg = gamma;
  float *part = nullptr, *prevq = nullptr, *dobj = obj;
  if (where == GNNUDS_MEM_HOST) {
    auto up = [&](int slot, const float** d, const float* src, size_t n) {
      float* q = nullptr;
      int r = grow(h, slot, n, &q);
      if (r) return r;
      *d = q;
      return cudaMemcpyAsync(q, src, n * sizeof(float), cudaMemcpyHostToDevice, h->stream) != cudaSuccess ? fail(h, GNNUDS_E_CUDA, "evaluate: copy failed") : 0;
    };
    if ((rc = up(11, &dx0, x0, nx0)) || (rc = up(12, &dex0, ex0, nex0)) || (rc = up(13, &db, b, nbb)) || (rc = up(14, &ddec, dec, ndec)) ||
        (rc = up(4, &dfw, flood_w, N)) || (rc = up(5, &dww, wwtp_w, N))) return rc;
    if (rough_w && (rc = up(6, &drw, rough_w, N))) return rc;
    if (gamma && (rc = up(7, &dg, gamma, (size_t)Th))) return rc;
  }
  if ((rc = grow(h, 15, (size_t)Bc * 128, &part)) || (rc = grow(h, 16, (size_t)Bc * N, &prevq))) return rc;
  if (obj_where == GNNUDS_MEM_HOST && (rc = grow(h, 17, (size_t)B, &dobj))) return rc;
  begin_timing(h);
  if (h->profiling) CU(cudaEventRecord(h->ev_a, h->stream));
  // the MPC / MBRL loops of the reference run predict_tf (mpc.py:250,575; mbrl.py:337)
  const int saved_variant = h->pp_variant;
  h->pp_variant = 1;
  struct Restore { gnnuds_handle* h; int v; ~Restore() { h->pp_variant = v; } } restore{h, saved_variant};
  for (int64_t b0 = 0; b0 < B; b0 += Bc) {
    const int64_t nb = std::min(Bc, B - b0);
    if ((rc = side_prepare(h, nb))) return rc;            // may allocate: before the capture
    char key[320];
    snprintf(key, sizeof key, "eval|%lld|%lld|%d|%d|%d|%d|%d|%d|%p|%p|%p|%p|%p|%p|%p|%p|%p", (long long)nb, (long long)b0, n_b, n_step, r_step, n_chunks,
             rule, h->side_on ? 1 : 0, (const void*)dx0, (const void*)dex0, (const void*)db, (const void*)ddec, (const void*)dfw, (const void*)dww,
             (const void*)drw, (const void*)dg, (void*)dobj);
    // the whole block - broadcasts, n_chunks forwards, feedback, objective - is one cached CUDA graph (a small population
    // is launch-latency-bound: ~45 kernels per chunk)
    auto block_body = [&]() -> int {
    // one state broadcast over the candidates of the chunk (mpc.py:237-240); q_in of the last observed step
      bcast_time_kernel<<<grid_for(h, nb * (long long)nx0, 256, 16), 256, 0, h->stream>>>(dx0, h->ro_x, nb, 0, 1, Ti, 0, Ti, (long long)N * c.n_in);
      bcast_time_kernel<<<grid_for(h, nb * (long long)nex0, 256, 16), 256, 0, h->stream>>>(dex0, h->ro_ex, nb, 0, 1, Ti, 0, Ti, (long long)M * c.e_in);
      init_prevq_kernel<<<grid_for(h, nb * N, 256, 16), 256, 0, h->stream>>>(dx0 + (size_t)(Ti - 1) * N * c.n_in, prevq, nb, N, c.n_in);
      h->launches += 3;
      for (int k = 0; k < n_chunks; ++k) {
        bcast_time_kernel<<<grid_for(h, nb * T * (long long)N * c.b_in, 256, 16), 256, 0, h->stream>>>(db, h->ro_b, nb, b0, n_b, (int)Th, k * T, T, (long long)N * c.b_in);
        dec_slice_kernel<<<grid_for(h, nb * T * c.n_act, 256, 16), 256, 0, h->stream>>>(ddec, h->ro_a, nb, b0, n_step, r_step, k * T, T, c.n_act);
        h->launches += 2;
        if ((rc = forward_any(h, h->ro_x, h->ro_b, h->ro_a, h->ro_ex, nb, h->out_y, h->out_ey, false))) return rc;
        ObjAccP op{h->out_y, nyc, prevq, dfw, dww, drw, dg, part, nb, T, N, k * T, k == 0 ? 1 : 0};
        obj_acc_kernel<<<grid_for(h, nb, 1, 16), 128, 0, h->stream>>>(op);
        ++h->launches;
        if (k + 1 < n_chunks) {
          FeedP p{};
          p.y = h->out_y; p.ny = nyc; p.ey = h->out_ey; p.e_out = c.e_out; p.b = h->ro_b; p.b_in = c.b_in;
          p.a = h->ro_a; p.n_act = c.n_act; p.act_of_edge = h->d_act_of_edge;
          p.x = h->ro_x; p.n_in = c.n_in; p.ex = h->ro_ex; p.e_in = c.e_in;
          p.S = nb * T; p.N = N; p.M = M; p.if_flood = 1; p.rule = rule;
          feedback_kernel<<<grid_for(h, p.S * (N + M), 256, 16), 256, 0, h->stream>>>(p);
          ++h->launches;
        }
        CU(cudaGetLastError());
      }
      obj_final_kernel<<<grid_for(h, nb, 1, 16), 128, 0, h->stream>>>(part, dobj + b0, nb);
      ++h->launches;
      return 0;
    };
    if ((rc = run_graphed(h, key, block_body))) return rc;
  }
  if (obj_where == GNNUDS_MEM_HOST) CU(cudaMemcpyAsync(obj, dobj, B * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
  if (h->profiling) CU(cudaEventRecord(h->ev_b, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  if (h->profiling) CU(cudaEventElapsedTime(&h->last_total_ms, h->ev_a, h->ev_b));
  return finish_timing(h);
}

// upload one raw event (replay / build_windows): device pointers of the four series
static int replay_upload(gnnuds_handle* h, const float* states, const float* perfs, const float* edge_states, const float* settings,
                         int64_t T, int32_t where, const float** ds, const float** dp, const float** de, const float** dse) {
  const auto& c = h->c;
  *ds = states; *dp = perfs; *de = edge_states; *dse = settings;
  if (where != GNNUDS_MEM_HOST) return 0;
  int rc;
  auto up = [&](int slot, const float** d, const float* src, size_t n) {
    float* q = nullptr;
    int r = grow(h, slot, n, &q);
    if (r) return r;
    *d = q;
    return cudaMemcpyAsync(q, src, n * sizeof(float), cudaMemcpyHostToDevice, h->stream) != cudaSuccess ? fail(h, GNNUDS_E_CUDA, "replay: copy failed") : 0;
  };
  if ((rc = up(18, ds, states, (size_t)T * c.n_node * 4)) || (rc = up(20, de, edge_states, (size_t)T * c.n_edge * c.e_in))) return rc;
  if (perfs && (rc = up(19, dp, perfs, (size_t)T * c.n_node))) return rc;
  if (settings && (rc = up(21, dse, settings, (size_t)T * c.n_act))) return rc;
  return 0;
}

static int replay_check(gnnuds_handle* h, const float* states, const float* perfs, const float* edge_states, const float* settings, int64_t T) {
  const auto& c = h->c;
  if (!states || !edge_states || T < 0) return fail(h, GNNUDS_E_INVALID, "replay: null series");
  if (c.if_flood && !perfs) return fail(h, GNNUDS_E_INVALID, "replay: an if_flood model needs the cumflooding series (main.py:339)");
  if (c.act && !settings) return fail(h, GNNUDS_E_INVALID, "replay: an act model needs the settings series (main.py:350)");
  if (c.n_in != 4 + (c.if_flood ? 1 : 0) || c.e_in != 4) return fail(h, GNNUDS_E_INVALID, "replay: node state [h,cuminflow,outflow,lateral] and 4 edge channels expected");
  if (c.b_in > 2) return fail(h, GNNUDS_E_INVALID, "replay: b_in %d", c.b_in);
  return 0;
}

static void replay_launch(gnnuds_handle* h, const float* ds, const float* dp, const float* de, const float* dse, int64_t T, int64_t w0,
                          int64_t nw, float* x, float* b, float* a, float* ex) {
  const auto& c = h->c;
  ReplayP p{ds, dp, de, dse, h->d_outfall, x, b, c.act ? a : nullptr, ex, w0, nw, (int)T, c.n_node, c.n_edge, c.seq_in, c.seq_out,
            std::max(c.seq_in, c.seq_out), c.n_in, c.b_in, c.e_in, c.n_act, c.if_flood ? 1 : 0};
  const long long work = nw * ((long long)c.seq_in * (c.n_node * c.n_in + c.n_edge * c.e_in) + (long long)c.seq_out * c.n_node * c.b_in);
  replay_windows_kernel<<<grid_for(h, work, 256, 16), 256, 0, h->stream>>>(p);
  ++h->launches;
}

int gnnuds_build_windows(gnnuds_handle* h, const float* states, const float* perfs, const float* edge_states, const float* settings,
                         int64_t T, int64_t w0, int64_t nw, float* x, float* b, float* a, float* ex, int32_t where) {
  int rc = check_ready(h);
  if (rc) return rc;
  const auto& c = h->c;
  if ((rc = replay_check(h, states, perfs, edge_states, settings, T))) return rc;
  if (!x || !b || !ex || (c.act && !a) || w0 < 0 || nw < 0 || w0 + nw > std::max<int64_t>(T - c.seq_out, 0)) return fail(h, GNNUDS_E_INVALID, "build_windows: bad window range");
  CU(cudaSetDevice(c.device));
  if (nw == 0) return 0;
  const float *ds, *dp, *de, *dse;
  if ((rc = replay_upload(h, states, perfs, edge_states, settings, T, where, &ds, &dp, &de, &dse))) return rc;
  const size_t nx = (size_t)nw * c.seq_in * c.n_node * c.n_in, nbb = (size_t)nw * c.seq_out * c.n_node * c.b_in,
               na = (size_t)nw * c.seq_out * std::max(c.n_act, 1), ne = (size_t)nw * c.seq_in * c.n_edge * c.e_in;
  if (where == GNNUDS_MEM_HOST) {
    float* tmp = nullptr;
    if ((rc = grow(h, 22, nx + nbb + na + ne, &tmp))) return rc;
    replay_launch(h, ds, dp, de, dse, T, w0, nw, tmp, tmp + nx, tmp + nx + nbb, tmp + nx + nbb + na);
    CU(cudaMemcpyAsync(x, tmp, nx * 4, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaMemcpyAsync(b, tmp + nx, nbb * 4, cudaMemcpyDeviceToHost, h->stream));
    if (c.act) CU(cudaMemcpyAsync(a, tmp + nx + nbb, (size_t)nw * c.seq_out * c.n_act * 4, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaMemcpyAsync(ex, tmp + nx + nbb + na, ne * 4, cudaMemcpyDeviceToHost, h->stream));
  } else {
    replay_launch(h, ds, dp, de, dse, T, w0, nw, x, b, a, ex);
  }
  CU(cudaGetLastError());
  CU(cudaStreamSynchronize(h->stream));
  return 0;
}

int gnnuds_replay(gnnuds_handle* h, const float* states, const float* perfs, const float* edge_states, const float* settings,
                  int64_t T, float* y, float* ey, int32_t where) {
  int rc = check_ready(h);
  if (rc) return rc;
  const auto& c = h->c;
  if ((rc = replay_check(h, states, perfs, edge_states, settings, T))) return rc;
  if (!y || !ey) return fail(h, GNNUDS_E_INVALID, "replay: null output");
  CU(cudaSetDevice(c.device));
  const int64_t W = T - c.seq_out;
  if (W <= 0) return 0;
  const float *ds, *dp, *de, *dse;
  if ((rc = replay_upload(h, states, perfs, edge_states, settings, T, where, &ds, &dp, &de, &dse))) return rc;
  const int64_t Bc = pick_chunk(h, W);
  if ((rc = ensure_workspace(h, Bc))) return rc;
  const int N = c.n_node, M = c.n_edge, To = c.seq_out;
  const int64_t sy = (int64_t)To * N * h->ny(), sey = (int64_t)To * M * c.e_out;
  const bool host = where == GNNUDS_MEM_HOST;
  if (host && (rc = ensure_pinned(h, 0, (size_t)(Bc * (sy + sey))))) return rc;
  begin_timing(h);
  if (h->profiling) CU(cudaEventRecord(h->ev_a, h->stream));
  float* sout[2][2] = {{h->out_y, h->out_ey}, {h->out2_y, h->out2_ey}};
  int64_t i = 0, prev_w0 = 0, prev_nw = 0;
  auto drain = [&](int k, int64_t w0, int64_t nw) -> int {
    CU(cudaEventSynchronize(h->ev_out[k]));
    parallel_copy({{y + w0 * sy, h->pin_out[k], (size_t)(nw * sy) * sizeof(float)},
                   {ey + w0 * sey, h->pin_out[k] + nw * sy, (size_t)(nw * sey) * sizeof(float)}});
    return 0;
  };
  for (int64_t w0 = 0; w0 < W; w0 += Bc, ++i) {
    const int64_t nw = std::min(Bc, W - w0);
    const int k = (int)(i & 1);
    replay_launch(h, ds, dp, de, dse, T, w0, nw, h->in_x, h->in_b, h->in_a, h->in_ex);
    float *oy = host ? sout[k][0] : y + w0 * sy, *oey = host ? sout[k][1] : ey + w0 * sey;
    if (host && i >= 2) CU(cudaStreamWaitEvent(h->stream, h->ev_out[k], 0));   // chunk i-2's results have left this output set
    if ((rc = forward_any(h, h->in_x, h->in_b, c.act ? h->in_a : nullptr, h->in_ex, nw, oy, oey, false))) return rc;
    if (host) {
      CU(cudaEventRecord(h->ev_comp[k], h->stream));
      CU(cudaStreamWaitEvent(h->s_d2h, h->ev_comp[k], 0));
      CU(cudaMemcpyAsync(h->pin_out[k], sout[k][0], nw * sy * sizeof(float), cudaMemcpyDeviceToHost, h->s_d2h));
      CU(cudaMemcpyAsync(h->pin_out[k] + nw * sy, sout[k][1], nw * sey * sizeof(float), cudaMemcpyDeviceToHost, h->s_d2h));
      CU(cudaEventRecord(h->ev_out[k], h->s_d2h));
      if (i >= 1 && (rc = drain(k ^ 1, prev_w0, prev_nw))) return rc;
      prev_w0 = w0; prev_nw = nw;
    }
  }
  if (host && i >= 1 && (rc = drain((int)((i - 1) & 1), prev_w0, prev_nw))) return rc;
  if (h->profiling) CU(cudaEventRecord(h->ev_b, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  if (h->profiling) CU(cudaEventElapsedTime(&h->last_total_ms, h->ev_a, h->ev_b));
  return finish_timing(h);
}

// Several recorded events in ONE call (the reference's test loop walks its events one by one, main.py:306-382): the
// windows of all events form one sequence that is cut into workspace chunks regardless of event boundaries - a
// 600-window event alone leaves the last launch of every persistent kernel a fraction of a wave - each chunk's windows
// are built on the device from the raw series of the events it touches, and the predictions of chunk c leave through
// the pinned ring while chunk c+1 computes.  Bitwise the results of gnnuds_replay event by event (chunking is
// bit-invisible).
int gnnuds_replay_batch(gnnuds_handle* h, int32_t n_events, const float* const* states, const float* const* perfs,
                        const float* const* edge_states, const float* const* settings, const int64_t* T, float* const* y,
                        float* const* ey) {
  int rc = check_ready(h);
  if (rc) return rc;
  const auto& c = h->c;
  if (n_events < 0 || !states || !edge_states || !T || !y || !ey) return fail(h, GNNUDS_E_INVALID, "replay_batch: null argument");
  for (int i = 0; i < n_events; ++i) {
    if ((rc = replay_check(h, states[i], perfs ? perfs[i] : nullptr, edge_states[i], settings ? settings[i] : nullptr, T[i]))) return rc;
    if (!y[i] || !ey[i]) return fail(h, GNNUDS_E_INVALID, "replay_batch: null output of event %d", i);
  }
  CU(cudaSetDevice(c.device));
  if (n_events == 0) return 0;
  const int N = c.n_node, M = c.n_edge, To = c.seq_out, Ti = c.seq_in;
  const int64_t sy = (int64_t)To * N * h->ny(), sey = (int64_t)To * M * c.e_out;
  const int64_t sx = (int64_t)Ti * N * c.n_in, sb = (int64_t)To * N * c.b_in, sa = (int64_t)To * c.n_act, se = (int64_t)Ti * M * c.e_in;
  int64_t Wmax = 0, Wsum = 0;
  for (int i = 0; i < n_events; ++i) { const int64_t W = std::max<int64_t>(T[i] - To, 0); Wmax = std::max(Wmax, W); Wsum += W; }
  if (Wsum == 0) return 0;
  const int64_t Bc = pick_chunk(h, Wsum);
  (void)Wmax;
  if ((rc = ensure_workspace(h, Bc)) || (rc = ensure_pinned(h, 0, (size_t)(Bc * (sy + sey))))) return rc;
  // the windows of all events form ONE sequence, cut into chunks of Bc windows regardless of event boundaries: a chunk
  // is a list of pieces (event, first window, count)
  struct Piece { int ev; int64_t w0, n; };
  std::vector<std::vector<Piece>> chunks(1);
  {
    int64_t fill = 0;
    for (int i = 0; i < n_events; ++i) {
      int64_t W = std::max<int64_t>(T[i] - To, 0), w0 = 0;
      while (W > 0) {
        if (fill == Bc) { chunks.emplace_back(); fill = 0; }
        const int64_t n = std::min(W, Bc - fill);
        chunks.back().push_back({i, w0, n});
        w0 += n; W -= n; fill += n;
      }
    }
  }
  // raw series of the events a chunk touches, two staging sets
  int64_t Tc_max = 0;
  for (auto& ch : chunks) { int64_t t = 0; for (auto& pc : ch) t += T[pc.ev]; Tc_max = std::max(Tc_max, t); }
  float* raw[2][4] = {{nullptr, nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr, nullptr}};
  for (int k = 0; k < 2; ++k) {
    if ((rc = grow(h, 18 + (k ? 24 : 0), (size_t)Tc_max * N * 4, &raw[k][0]))) return rc;                         // slots 18 / 42
    if ((rc = grow(h, 19 + (k ? 24 : 0), (size_t)Tc_max * N, &raw[k][1]))) return rc;                            // 19 / 43
    if ((rc = grow(h, 20 + (k ? 24 : 0), (size_t)Tc_max * M * c.e_in, &raw[k][2]))) return rc;                   // 20 / 44
    if ((rc = grow(h, 21 + (k ? 24 : 0), (size_t)Tc_max * std::max(c.n_act, 1), &raw[k][3]))) return rc;         // 21 / 45
  }
  begin_timing(h);
  if (h->profiling) CU(cudaEventRecord(h->ev_a, h->stream));
  float* sout[2][2] = {{h->out_y, h->out_ey}, {h->out2_y, h->out2_ey}};
  auto drain = [&](int k, const std::vector<Piece>& ch) -> int {
    CU(cudaEventSynchronize(h->ev_out[k]));
    std::vector<CopyJob> jobs;
    int64_t w = 0, nwc = 0;
    for (auto& pc : ch) nwc += pc.n;
    for (auto& pc : ch) {
      jobs.push_back({y[pc.ev] + pc.w0 * sy, h->pin_out[k] + w * sy, (size_t)(pc.n * sy) * sizeof(float)});
      jobs.push_back({ey[pc.ev] + pc.w0 * sey, h->pin_out[k] + nwc * sy + w * sey, (size_t)(pc.n * sey) * sizeof(float)});
      w += pc.n;
    }
    parallel_copy(jobs);
    return 0;
  };
  for (size_t ci = 0; ci < chunks.size(); ++ci) {
    const auto& ch = chunks[ci];
    const int k = (int)(ci & 1);
    // raw set k was last read by the window builders of chunk ci-2; the copies below are on the compute stream, ordered
    // after them
    int64_t toff = 0, woff = 0;
    for (auto& pc : ch) {
      const int i = pc.ev;
      const int64_t Te = T[i];
      float *ds = raw[k][0] + toff * N * 4, *dp = raw[k][1] + toff * N, *de = raw[k][2] + toff * M * c.e_in, *dse = raw[k][3] + toff * c.n_act;
      CU(cudaMemcpyAsync(ds, states[i], (size_t)Te * N * 4 * sizeof(float), cudaMemcpyHostToDevice, h->stream));
      CU(cudaMemcpyAsync(de, edge_states[i], (size_t)Te * M * c.e_in * sizeof(float), cudaMemcpyHostToDevice, h->stream));
      if (c.if_flood) CU(cudaMemcpyAsync(dp, perfs[i], (size_t)Te * N * sizeof(float), cudaMemcpyHostToDevice, h->stream));
      if (c.act) CU(cudaMemcpyAsync(dse, settings[i], (size_t)Te * c.n_act * sizeof(float), cudaMemcpyHostToDevice, h->stream));
      // windows pc.w0 .. pc.w0+pc.n-1 of event i land at rows woff.. of the chunk (the builder writes window w at index w - w0)
      replay_launch(h, ds, c.if_flood ? dp : nullptr, de, c.act ? dse : nullptr, Te, pc.w0, pc.n, h->in_x + woff * sx, h->in_b + woff * sb,
                    h->in_a + woff * sa, h->in_ex + woff * se);
      toff += Te; woff += pc.n;
    }
    const int64_t nw = woff;
    if (nw) {
      if (ci >= 2) CU(cudaStreamWaitEvent(h->stream, h->ev_out[k], 0));     // chunk ci-2's results have left this output set
      if ((rc = forward_any(h, h->in_x, h->in_b, c.act ? h->in_a : nullptr, h->in_ex, nw, sout[k][0], sout[k][1], false))) return rc;
    }
    CU(cudaEventRecord(h->ev_comp[k], h->stream));
    CU(cudaStreamWaitEvent(h->s_d2h, h->ev_comp[k], 0));
    if (nw) {
      CU(cudaMemcpyAsync(h->pin_out[k], sout[k][0], nw * sy * sizeof(float), cudaMemcpyDeviceToHost, h->s_d2h));
      CU(cudaMemcpyAsync(h->pin_out[k] + nw * sy, sout[k][1], nw * sey * sizeof(float), cudaMemcpyDeviceToHost, h->s_d2h));
    }
    CU(cudaEventRecord(h->ev_out[k], h->s_d2h));
    if (ci >= 1 && (rc = drain(k ^ 1, chunks[ci - 1]))) return rc;
  }
  if ((rc = drain((int)((chunks.size() - 1) & 1), chunks.back()))) return rc;
  if (h->profiling) CU(cudaEventRecord(h->ev_b, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  if (h->profiling) CU(cudaEventElapsedTime(&h->last_total_ms, h->ev_a, h->ev_b));
  return finish_timing(h);
}

int gnnuds_alloc_pinned(void** p, int64_t bytes) {
  if (!p) return GNNUDS_E_INVALID;
  return cudaMallocHost(p, (size_t)bytes) == cudaSuccess ? 0 : GNNUDS_E_NOMEM;
}
int gnnuds_free_pinned(void* p) { return cudaFreeHost(p) == cudaSuccess ? 0 : GNNUDS_E_CUDA; }
int gnnuds_alloc_device(gnnuds_handle* h, void** p, int64_t bytes) {
  if (!h || !p) return GNNUDS_E_INVALID;
  CU(cudaSetDevice(h->c.device));
  return cudaMalloc(p, (size_t)bytes) == cudaSuccess ? 0 : fail(h, GNNUDS_E_NOMEM, "cudaMalloc(%lld) failed", (long long)bytes);
}
int gnnuds_free_device(gnnuds_handle* h, void* p) {
  if (!h) return GNNUDS_E_INVALID;
  CU(cudaSetDevice(h->c.device));
  CU(cudaFree(p));
  return 0;
}
int gnnuds_memcpy(gnnuds_handle* h, void* dst, const void* src, int64_t bytes, int32_t kind) {
  if (!h) return GNNUDS_E_INVALID;
  CU(cudaSetDevice(h->c.device));
  CU(cudaMemcpyAsync(dst, src, (size_t)bytes, kind == 1 ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return 0;
}
int gnnuds_synchronize(gnnuds_handle* h) {
  if (!h) return GNNUDS_E_INVALID;
  CU(cudaStreamSynchronize(h->stream));
  return 0;
}
int64_t gnnuds_launch_count(const gnnuds_handle* h) { return h ? h->launches : 0; }
int gnnuds_last_timing(const gnnuds_handle* h, float* total_ms, float* gemm_ms) {
  if (!h) return GNNUDS_E_INVALID;
  if (total_ms) *total_ms = h->last_total_ms;
  if (gemm_ms) *gemm_ms = h->last_gemm_ms;
  return 0;
}
int gnnuds_kernel_timing(const gnnuds_handle* h, int32_t family, float* ms, int64_t* launches) {
  if (!h || family < 0 || family > 3) return GNNUDS_E_INVALID;
  if (ms) *ms = h->fam_ms[family];
  if (launches) *launches = h->fam_launches[family];
  return 0;
}
int gnnuds_set_profiling(gnnuds_handle* h, int32_t on) {
  if (!h) return GNNUDS_E_INVALID;
  h->profiling = on;
  return 0;
}
int gnnuds_debug_stop_after(gnnuds_handle* h, int32_t stage) {
  if (!h) return GNNUDS_E_INVALID;
  h->debug_stop = stage;
  return 0;
}
int gnnuds_debug_umma_selftest(gnnuds_handle* h, int32_t mode, const float* act, int32_t Ra, int32_t Ca, const float* mat,
                               int32_t Rm, int32_t Km, int32_t shift, int32_t swap, float* out) {
  if (!h || !act || !mat || !out) return GNNUDS_E_INVALID;
  CU(cudaSetDevice(h->c.device));
  const int Nout = mode == 0 ? Rm : Ca;
  if (Ca % 16 || Km % 16 || Nout % 16 || Nout > 256 || Nout < 16) return fail(h, GNNUDS_E_INVALID, "selftest: bad shape");
  if (mode == 0 && (Km != Ca || Ra < shift + 128 || shift < 0)) return fail(h, GNNUDS_E_INVALID, "selftest mode 0: bad shape");
  if (mode == 1 && (Rm != 128 || Ra < Km)) return fail(h, GNNUDS_E_INVALID, "selftest mode 1: bad shape");
  const int ps_img = Rm * 16;
  std::vector<uint16_t> img = pack_kmajor(mat, Rm, Km, Km, 1, Rm, Km, ps_img);
  float *d_act = nullptr, *d_out = nullptr;
  uint16_t* d_img = nullptr;
  int rc = 0;
  if ((rc = upload(h, &d_act, act, (size_t)Ra * Ca))) return rc;
  if ((rc = upload(h, &d_img, img.data(), img.size()))) { cudaFree(d_act); return rc; }
  if (cudaMalloc(&d_out, (size_t)128 * Nout * sizeof(float)) != cudaSuccess) { cudaFree(d_act); cudaFree(d_img); return fail(h, GNNUDS_E_NOMEM, "selftest"); }
  SelfTestP p{d_act, d_img, d_out, Ra, Ca, Rm, Km, shift, swap, mode, Nout};
  const size_t smem = (((size_t)(Ca / 8) * (Ra + 1) * 16 + 127) & ~(size_t)127) + (size_t)(Km / 8) * ps_img + 256;
  cudaError_t e = cudaFuncSetAttribute(umma_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e == cudaSuccess) {
    umma_selftest_kernel<<<1, 256, smem, h->stream>>>(p);
    ++h->launches;
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
  if (e == cudaSuccess) e = cudaMemcpy(out, d_out, (size_t)128 * Nout * sizeof(float), cudaMemcpyDeviceToHost);
  cudaFree(d_act); cudaFree(d_img); cudaFree(d_out);
  if (e != cudaSuccess) return fail(h, GNNUDS_E_CUDA, "selftest: %s", cudaGetErrorString(e));
  return 0;
}

int gnnuds_debug_fetch(gnnuds_handle* h, int32_t which, float* out, int64_t max_floats, int64_t* rows, int32_t* width) {
  if (!h || which < 0 || which > 1 || !h->dbg_ptr[which]) return GNNUDS_E_INVALID;
  CU(cudaSetDevice(h->c.device));
  const int64_t n = h->dbg_rows[which] * h->dbg_width[which];
  if (rows) *rows = h->dbg_rows[which];
  if (width) *width = h->dbg_width[which];
  if (out) {
    if (n > max_floats) return fail(h, GNNUDS_E_INVALID, "debug_fetch: buffer too small");
    CU(cudaStreamSynchronize(h->stream));
    if (h->tc()) {
      std::vector<uint16_t> tmp((size_t)n);
      CU(cudaMemcpy(tmp.data(), h->dbg_ptr[which], n * 2, cudaMemcpyDeviceToHost));
      for (int64_t i = 0; i < n; ++i) out[i] = bf2f(tmp[i]);
    } else {
      CU(cudaMemcpy(out, h->dbg_ptr[which], n * sizeof(float), cudaMemcpyDeviceToHost));
    }
  }
  return 0;
}

}  // extern "C"

// =====================================================================================================================
// gnnuds_objective_static: the astlingen objective with NO allocation (include/gnnuds.h).  No counterpart in the
// reference; added for callers that cannot call cudaMalloc (CUDA graph capture,
// tight serving loops, harnesses that forbid it).
//
// The chunked objective (obj_acc_kernel, kernels.cuh) carries two pieces of scratch from one time slab to the next: the
// 128 per-thread partial sums of every candidate and q_in of the slab's last step.  gnnuds_evaluate keeps them in
// grow-only buffers of the handle (cudaMalloc on first use); here they live in ONE file-scope device array of a fixed
// size that every GPU which loads the library holds until the process exits.  All calls (all handles, all streams) on a
// device share it: calls that overlap on the GPU race.  Nothing in it has to start at zero - the first slab overwrites
// the partial sums and q_in - so a call is not affected by what the previous one left behind.
// =====================================================================================================================
namespace {

static __device__ __align__(256) float g_obj_scratch[GNNUDS_STATIC_SCRATCH_BYTES / sizeof(float)];

// In host code the array's name is only a handle for the runtime, not a device address: the address a kernel can use
// comes from cudaGetSymbolAddress, once per device, with that device current.
constexpr int GU_STATIC_MAX_DEV = 64;
float* g_obj_scratch_ptr[GU_STATIC_MAX_DEV] = {};
std::mutex g_obj_scratch_mu;

// time slab t0..t1-1 of the FULL prediction tensor y (B,T,N,ny).  Same ownership as obj_acc_kernel / objective_kernel:
// thread tid of candidate i adds the terms g = t*N + n with g % 128 == tid in increasing g, so the final sum is bitwise
// the one objective_kernel returns.
struct ObjSlabP {
  const float* y; int ny;
  const float* q_in0;
  const float* flood_w; const float* wwtp_w; const float* rough_w; const float* gamma;
  float* part;      // scratch (B,128): running partial sums
  float* prev_q;    // scratch (B,N): q_in of step t0-1, written by the slab before
  long long B; int T, N, t0, t1;
};
__global__ void __launch_bounds__(128) obj_slab_static_kernel(ObjSlabP p) {
  for (long long i = blockIdx.x; i < p.B; i += gridDim.x) {
    const int tid = threadIdx.x;
    float acc = p.t0 == 0 ? 0.f : p.part[i * 128 + tid];
    const int g0 = p.t0 * p.N;
    for (int g = g0 + ((tid - g0) % 128 + 128) % 128; g < p.t1 * p.N; g += 128) {
      const int t = g / p.N, n = g % p.N;
      const float* ys = p.y + ((i * p.T + t) * p.N + n) * p.ny;
      const float gm = p.gamma ? p.gamma[t] : 1.f;
      float v = ys[p.ny - 1] * p.flood_w[n] + ys[1] * p.wwtp_w[n];
      if (p.rough_w) {
        const float prev = t > p.t0 ? (ys - (long long)p.N * p.ny)[1] : (p.t0 == 0 ? p.q_in0[i * p.N + n] : p.prev_q[i * p.N + n]);
        v += fabsf(ys[1] - prev) * p.rough_w[n];
      }
      acc += v * gm;
    }
    p.part[i * 128 + tid] = acc;
    __syncthreads();                                   // every read of prev_q of this candidate is done
    for (int n = tid; n < p.N; n += 128) p.prev_q[i * p.N + n] = p.y[((i * p.T + p.t1 - 1) * p.N + n) * p.ny + 1];
  }
}

}  // namespace

extern "C" int gnnuds_objective_static(gnnuds_handle* h, const float* y, const float* q_in0, int64_t B, int32_t T,
                                       const float* flood_w, const float* wwtp_w, const float* rough_w, const float* gamma,
                                       float* obj, int32_t sync) {
  if (!h) return GNNUDS_E_INVALID;
  if (!y || !q_in0 || !flood_w || !wwtp_w || !obj || B < 0 || T < 1) return fail(h, GNNUDS_E_INVALID, "objective_static: bad argument");
  const auto& c = h->c;
  const int N = c.n_node, nyc = h->ny();
  const long long need = (long long)B * (128 + N), have = (long long)(GNNUDS_STATIC_SCRATCH_BYTES / sizeof(float));
  if (need > have)
    return fail(h, GNNUDS_E_INVALID,
                "objective_static: B * (128 + n_node) = %lld x %d = %lld floats of scratch, the static array holds %lld "
                "(GNNUDS_STATIC_SCRATCH_BYTES = %d); split the population or use gnnuds_objective",
                (long long)B, 128 + N, need, have, (int)GNNUDS_STATIC_SCRATCH_BYTES);
  if (c.device < 0 || c.device >= GU_STATIC_MAX_DEV) return fail(h, GNNUDS_E_INVALID, "objective_static: device %d", c.device);
  CU(cudaSetDevice(c.device));
  if (B == 0) return 0;
  float* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_obj_scratch_mu);
    if (!g_obj_scratch_ptr[c.device]) {
      float* ptr = nullptr;
      cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_obj_scratch);
      if (e != cudaSuccess || !ptr) return fail(h, GNNUDS_E_CUDA, "objective_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      g_obj_scratch_ptr[c.device] = ptr;
    }
    scratch = g_obj_scratch_ptr[c.device];
  }
  ObjSlabP p{y, nyc, q_in0, flood_w, wwtp_w, rough_w, gamma, scratch, scratch + (size_t)B * 128, (long long)B, T, N, 0, 0};
  const int grid = grid_for(h, B, 1, 16);
  const int cut = T >= 2 ? T / 2 : T;                  // two slabs: the scratch really is the carried state
  for (int t0 = 0; t0 < T; t0 = p.t1) {
    p.t0 = t0; p.t1 = t0 == 0 ? cut : T;
    obj_slab_static_kernel<<<grid, 128, 0, h->stream>>>(p);
    ++h->launches;
  }
  obj_final_kernel<<<grid, 128, 0, h->stream>>>(scratch, obj, B);
  ++h->launches;
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess && sync) e = cudaStreamSynchronize(h->stream);
  if (e != cudaSuccess) return fail(h, GNNUDS_E_CUDA, "objective_static: %s", cudaGetErrorString(e));
  return 0;
}

#include "backward_host.inc"
