/*
 * gnnuds.h - C ABI of the B200-native GNN-UDS Emulator rollout (libgnnuds.so).
 *
 * The reference (Zhiyu014/GNN-UDS) has no FFI; its boundary for this path is the
 * Python class surrogate/emulator.py::Emulator.  Each entry point below names the
 * reference interface it replaces (file:line under /root/reference/surrogate).
 * The Python shim gnn_uds_b200/emulator.py binds these through ctypes
 * (see INTEGRATION.md).
 *
 * Conventions: every call returns 0 on success, a negative GNNUDS_E_* code
 * otherwise; nothing throws across the boundary; the library never takes
 * ownership of caller memory; one handle = one CUDA device + one stream, not
 * thread-safe; handles must not cross processes.  All arrays are C-contiguous,
 * little-endian float32 / int32 unless stated.
 */
#ifndef GNNUDS_H
#define GNNUDS_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GNNUDS_ABI_VERSION 2

enum {
  GNNUDS_OK = 0,
  GNNUDS_E_INVALID = -1,   /* bad argument / unsupported variant */
  GNNUDS_E_STATE = -2,     /* call order (graph/weights/norm not set) */
  GNNUDS_E_CUDA = -3,      /* CUDA runtime error, see gnnuds_last_error */
  GNNUDS_E_NOMEM = -4
};

enum { GNNUDS_CONV_NONE = 0 /* NN-cat: conv falsy, emulator.py:180,231-233 */, GNNUDS_CONV_GCN = 1, GNNUDS_CONV_GAT = 2 };  /* emulator.py:129-152 */
enum { GNNUDS_REC_CONV1D = 1, GNNUDS_REC_GRU = 2, GNNUDS_REC_LSTM = 3 }; /* emulator.py:154-163 */
enum { GNNUDS_ACT_LINEAR = 0, GNNUDS_ACT_RELU = 1, GNNUDS_ACT_TANH = 2, GNNUDS_ACT_SIGMOID = 3 };
enum { GNNUDS_PREC_FP32 = 0,    /* fp32 SIMT, parity mode (<=1e-5 rel per step) */
       GNNUDS_PREC_BF16_TC = 1  /* 16-bit operands on tcgen05 (fp16 by default, bf16 with -DGNNUDS_TC_BF16; see
                               gnnuds_tc_operand_is_half), fp32 accumulate */ };
enum { GNNUDS_MEM_HOST = 0, GNNUDS_MEM_DEVICE = 1 };
/* feedback rule of the autoregressive chunk loop */
enum { GNNUDS_ROLL_MPC = 1,     /* mpc.py:565-582 */
       GNNUDS_ROLL_MBRL = 2,    /* mbrl.py:337-346 */
       GNNUDS_ROLL_NORM = 3     /* emulator.py:400-438 (normalised space) */ };

/* POD mirror of the attributes Emulator.__init__ reads off `args` (emulator.py:48-127). */
typedef struct gnnuds_config {
  int32_t abi_version;     /* GNNUDS_ABI_VERSION */
  int32_t n_node, n_edge;  /* state_shape[0], edge_state_shape[0] */
  int32_t n_in;            /* node input channels incl. flood flag (emulator.py:49,69-70) */
  int32_t b_in;            /* 1, or 2 with tide (emulator.py:52) */
  int32_t e_in;            /* edge input channels (4) */
  int32_t n_out;           /* node head width after the edge_fusion cut (emulator.py:56,85) */
  int32_t e_out;           /* e_in - 1 */
  int32_t seq_in, seq_out;
  int32_t embed, hidden, kernel_size, n_sp_layer, n_tp_layer;
  int32_t conv, recurrent, activation;
  int32_t resnet, if_flood, edge_fusion, act, graph_base, tide;
  int32_t n_act;           /* columns of the settings tensor `a` */
  float   epsilon;         /* emulator.py:72, flood gating rule in constrain */
  int32_t precision;       /* GNNUDS_PREC_* */
  int32_t device;          /* CUDA ordinal */
  int32_t max_chunk;       /* scenarios per internal pass; 0 = auto from workspace_mb */
  int32_t workspace_mb;    /* activation workspace budget in MB; 0 = 8192 (the tensor-core path fits twice the scenarios in it) */
} gnnuds_config;

/* per-network vectors, get_args (envs/scenario/base.py:277-332) + emulator.py:71-98 */
typedef struct gnnuds_graph {
  /* node adjacency filter in CSR (joint node+edge graph when graph_base != 0).
     GAT: pattern of (adj>0) with forced diagonal, val = NULL.
     GCN: val = gcn_filter(adj) (emulator.py:133, spektral.utils.gcn_filter). */
  const int32_t* row_ptr; const int32_t* col_idx; const float* val; int32_t n; int32_t nnz;
  /* edge (line-graph) adjacency filter, unused when graph_base != 0 */
  const int32_t* e_row_ptr; const int32_t* e_col_idx; const float* e_val; int32_t m; int32_t e_nnz;
  const int32_t* edges;        /* (n_edge,2) from,to  - base.py:349-365 */
  const int32_t* act_of_edge;  /* (n_edge) column of `a` driving this edge or -1 - emulator.py:385-398 */
  const int32_t* act_out_of_node; /* (n_node) column of `a` or -1 - emulator.py:364-383 (a_out) */
  const int32_t* act_in_of_node;  /* (n_node) column of `a` or -1 (a_in) */
  const float* hmin; const float* hmax; const float* is_outfall;   /* (n_node) */
  const float* ehmax;                                               /* (n_edge) */
  /* pumped / offset networks (chaohu, hague): NULL = all zero.  emulator.py:590-598 (pumped-storage depth),
     649-653 (inlet offset gate), 656-667 (rated pump flow); envs/scenario/chaohu.py:203-226 */
  const float* pump;      /* (n_edge) rated pump flow of the link, 0 = not a pump */
  const float* offset;    /* (n_edge) inlet offset of the link */
  const float* pump_in;   /* (n_node) rated inflow by pumps (non-edge-fusion models) */
  const float* pump_out;  /* (n_node) rated outflow by pumps */
  const float* area;      /* (n_node) storage area (emulator.py:591,596) */
} gnnuds_graph;

typedef struct gnnuds_handle gnnuds_handle;

/* Emulator.__init__ + build_network (emulator.py:48-127,166-341). */
int gnnuds_create(const gnnuds_config* cfg, gnnuds_handle** out);
int gnnuds_destroy(gnnuds_handle* h);
const char* gnnuds_last_error(const gnnuds_handle* h);   /* h may be NULL: create() errors */

/* Graph tensors: Emulator.get_conv filters + vectors read in __init__. */
int gnnuds_set_graph(gnnuds_handle* h, const gnnuds_graph* g);

/* Number of weight tensors build_network creates, in creation order, and their sizes.
   sizes may be NULL to query the count. */
int gnnuds_weight_layout(gnnuds_handle* h, int32_t* n_tensors, int64_t* sizes);

/* Emulator.load -> model.load_weights (emulator.py:833-839): all tensors in Keras
   creation order and Keras layout, concatenated into one blob. offsets[i] is the
   float offset of tensor i, sizes[i] its element count (checked). */
int gnnuds_set_weights(gnnuds_handle* h, const float* blob, int64_t n_floats,
                       const int64_t* offsets, const int64_t* sizes, int32_t n_tensors);

/* Emulator.set_norm (emulator.py:794-800). norm_* are (2, n, c): [0]=max, [1]=min.
   c_y is the channel count of norm_y (>= channels denormalised). norm_b/norm_e as given. */
int gnnuds_set_norm(gnnuds_handle* h, const float* norm_x, int32_t c_x,
                    const float* norm_b, int32_t c_b,
                    const float* norm_y, int32_t c_y,
                    const float* norm_e, int32_t c_e);

/* use_adj (emulator.py:92,178-180,268-271,343-362; act models, GAT): the block-2 node convolution attends over a
   per-sample adjacency - the static one with the setting of every controlled link written into its (from,to) entry
   and cast to int (get_adj_action: `adj.astype(int)`), so an entry survives iff trunc(value) != 0.  act_entry (nnz of
   the node filter set by gnnuds_set_graph): 1 + column of `a` for the entry of a controlled link, 0 otherwise;
   base_keep (nnz): trunc(adjacency value) != 0 for the other entries.  The diagonal is always kept (spektral
   add_self_loops).  GCN + use_adj is rejected: the reference's call raises inside spektral (4-D gcn_filter). */
int gnnuds_set_adj_action(gnnuds_handle* h, const int32_t* act_entry, const uint8_t* base_keep);

/* Which of the reference's two post-processing variants gnnuds_predict / gnnuds_simulate follow:
   0 = post_proc (NumPy, emulator.py:643-678; what Emulator.predict and simulate run),
   1 = post_proc_tf (emulator.py:680-725; what predict_tf runs).  They differ on pumped networks only: the TF
   variant applies the rated pump flow to the link flows only when every link is a pump (pump.min()>0, :698) and
   gates pump_in / pump_out on h > 0 instead of h > 0.01 (:710-711).  gnnuds_rollout always uses variant 1
   (mpc.py:575, mbrl.py:337, emulator.py:413 call the TF variant). */
int gnnuds_set_post_proc_variant(gnnuds_handle* h, int32_t variant);

/* 16-bit operand type of this build's tensor-core path: 16 = IEEE fp16 (default), 0 = bfloat16 (-DGNNUDS_TC_BF16).
   No handle needed. */
int gnnuds_tc_operand_is_half(void);
/* fp32 FFMA peak of the handle's device in TFLOP/s, measured now with a register-only FMA kernel (the roofline
 * denominator of the fp32 parity path; SURVEY.md 8(d) asks for a measured figure instead of the nominal 74). */
int gnnuds_measure_simt_tflops(gnnuds_handle* h, float* tflops);

/* Channels of the node output y ([h,q_us,q_ds,(flood_prob),q_w]) and edge output. */
int gnnuds_output_dims(const gnnuds_handle* h, int32_t* y_channels, int32_t* ey_channels);

/* Emulator.predict / predict_tf (emulator.py:566-641): physical units in, physical out.
   x  (B, seq_in, n_node, n_in)   b (B, seq_out, n_node, b_in)
   a  (B, seq_out, n_act) or NULL ex (B, seq_in, n_edge, e_in)
   y  (B, seq_out, n_node, y_channels)   ey (B, seq_out, n_edge, e_out)
   where = GNNUDS_MEM_HOST (copies inside) or GNNUDS_MEM_DEVICE (pointers on cfg.device). */
int gnnuds_predict(gnnuds_handle* h, const float* x, const float* b, const float* a,
                   const float* ex, int64_t B, float* y, float* ey, int32_t where);

/* Emulator.simulate (emulator.py:521-564): T' teacher-forced windows evaluated as one batch.
   runoff (T', seq_r, n_node, b_in) with seq_r >= seq_out (first seq_out steps used). */
int gnnuds_simulate(gnnuds_handle* h, const float* states, const float* runoff, int32_t seq_r,
                    const float* a, const float* edge_states, int64_t n_windows,
                    float* y, float* ey, int32_t where);

/* Autoregressive chunk loop, n_chunks forwards of seq_out steps with the feedback rule
   `rule` applied on the device between chunks (mpc.py:565-582, mbrl.py:337-346,
   emulator.py:400-438). Requires seq_in == seq_out for MPC/MBRL as in the reference.
   x0 (B,seq_in,n_node,n_in)  b (B,n_chunks*seq_out,n_node,b_in)  a (B,n_chunks*seq_out,n_act)|NULL
   ex0 (B,seq_in,n_edge,e_in)  y (B,n_chunks*seq_out,n_node,y_ch)  ey (B,n_chunks*seq_out,n_edge,e_out) */
int gnnuds_rollout(gnnuds_handle* h, int32_t rule, const float* x0, const float* b, const float* a,
                   const float* ex0, int64_t B, int32_t n_chunks, float* y, float* ey, int32_t where);

/* astlingen.objective_pred_tf (envs/scenario/astlingen.py:75-99): per-candidate scalar.
   y (B,T,n_node,y_ch) as returned by predict/rollout; q_in0 (B,n_node) = state[:,-1,:,1].
   flood_w (n_node) weight of q_w per node (0 = not a target); wwtp_w (n_node) weight of
   q_in[1:]; rough_w (n_node) weight of |diff_t q_in| (pass NULL for rough=False);
   gamma (T) or NULL. obj (B). */
int gnnuds_objective(gnnuds_handle* h, const float* y, const float* q_in0, int64_t B, int32_t T,
                     const float* flood_w, const float* wwtp_w, const float* rough_w,
                     const float* gamma, float* obj, int32_t where);

/* The same objective with NO allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Every pointer is DEVICE memory on cfg.device, laid out as for gnnuds_objective (y (B,T,n_node,y_ch), q_in0 (B,n_node),
 * flood_w / wwtp_w / rough_w|NULL (n_node), gamma|NULL (T), obj (B)).  The objective is accumulated in two time slabs
 * whose carried state (128 partial sums and the last q_in per candidate) lives in a file-scope `static __device__` array
 * of GNNUDS_STATIC_SCRATCH_BYTES that every GPU which loads the library holds until exit; its device address is taken
 * once per device with cudaGetSymbolAddress.  The call makes no cudaMalloc / cudaFree / allocating library call.
 * Limit: B * (128 + n_node) * 4 bytes <= GNNUDS_STATIC_SCRATCH_BYTES (astlingen, 30 nodes: B <= 6636), else
 * GNNUDS_E_INVALID with the sizes in gnnuds_last_error.  All calls on a device share the array: calls that overlap on
 * the GPU (two handles, two host threads) race; calls in sequence are independent (nothing in the array has to start at
 * zero, the first slab overwrites it).  Runs on the handle's stream; sync != 0 waits for it, sync == 0 only enqueues
 * (gnnuds_synchronize waits).  Bitwise equal to gnnuds_objective. */
#define GNNUDS_STATIC_SCRATCH_BYTES (4 << 20)
int gnnuds_objective_static(gnnuds_handle* h, const float* y, const float* q_in0, int64_t B, int32_t T,
                            const float* flood_w, const float* wwtp_w, const float* rough_w,
                            const float* gamma, float* obj, int32_t sync);

/* MPC population evaluation as ONE device-resident call (mpc.py:210-259 pred_emu/pred_emu_tf, :565-582 chunked loop,
 * envs/scenario/astlingen.py:75-99 objective): B candidates share one observed state x0 (seq_in,N,n_in) / ex0
 * (seq_in,M,e_in) - the reference tiles them on the host (mpc.py:237-240) - and n_b runoff scenarios b
 * (n_b, n_chunks*seq_out, N, b_in); candidate i runs scenario i % n_b (n_b = `stochastic`, else 1).  dec is the
 * decision matrix (B, n_step, n_act); every setting is held r_step steps and the last one to the end of the horizon
 * (mpc.py:211-218).  n_chunks == 1 is one forward, n_chunks > 1 the autoregressive chunk loop with feedback `rule`
 * (GNNUDS_ROLL_MPC / GNNUDS_ROLL_MBRL).  Predictions never leave the device: the objective (weights per node as in
 * gnnuds_objective, gamma per horizon step or NULL) is accumulated chunk by chunk and only obj (B) is written, to host
 * (obj_where GNNUDS_MEM_HOST) or device memory.  `where` says where x0, ex0, b, dec and the weight vectors live.
 * Bitwise equal to gnnuds_predict / gnnuds_rollout followed by gnnuds_objective on the fp32 path. */
int gnnuds_evaluate(gnnuds_handle* h, int32_t rule, const float* x0, const float* ex0, const float* b, int32_t n_b,
                    const float* dec, int32_t n_step, int32_t r_step, int64_t B, int32_t n_chunks, const float* flood_w,
                    const float* wwtp_w, const float* rough_w, const float* gamma, float* obj, int32_t where,
                    int32_t obj_where);

/* Reverse pass of predict_tf w.r.t. its INPUTS (no weight gradients): the vector-Jacobian product that
 * tf.GradientTape evaluates in mpc_problem_gr (mpc.py:511-627: gradient_fn :596-607, pred_fit :512-548 differentiate
 * objective_pred_tf(predict_tf(state, runoff, settings, edge_state)) w.r.t. the settings).  x, b, a, ex as for
 * gnnuds_predict; gy (B,seq_out,n_node,y_channels) / gey (B,seq_out,n_edge,e_out) are the cotangents of the two outputs.
 * Written: y / ey (the forward results, NULL = not wanted), ga (B,seq_out,n_act), gx (B,seq_in,n_node,n_in),
 * gex (B,seq_in,n_edge,e_in) - each may be NULL.  The forward runs with post_proc_tf (variant 1) and keeps its
 * activations on a tape (grow-only arena of the handle); sub-gradients as TensorFlow defines them (relu'(0) = 0,
 * clip_by_value passes the gradient on the closed interval, comparisons carry none).  The flood-flag input channel and
 * thresholds (flood prob > 0.5, epsilon mask) are treated as constants, as GradientTape does.
 * Supported: conv models (GCN / GAT; per-side graphs, the joint graph of graph_base != 0, use_adj on the per-side graph)
 * and the NN-cat variant, Conv1D / GRU / LSTM temporal nets, pumped / offset networks (post_proc_tf rules, pumped-storage
 * scan), fp32 handles.  GNNUDS_E_INVALID for tensor-core handles and for graphs whose per-sample rows exceed the
 * aggregation VJP's shared memory (3 * rows * embed floats <= 227 KB). */
int gnnuds_predict_vjp(gnnuds_handle* h, const float* x, const float* b, const float* a, const float* ex, int64_t B,
                       const float* gy, const float* gey, float* y, float* ey, float* ga, float* gx, float* gex,
                       int32_t where);

/* mpc_problem_gr.gradient_fn (mpc.py:596-607) for the astlingen objective (envs/scenario/astlingen.py:75-99):
 * obj (B) and grad_a (B, n_chunks*seq_out, n_act) = d obj_i / d a_i for every candidate i.  n_chunks == 1 is one
 * predict_tf; n_chunks > 1 differentiates through the autoregressive chunk loop (mpc.py:565-582, feedback `rule`
 * GNNUDS_ROLL_MPC or GNNUDS_ROLL_MBRL): the forward sweep keeps the tapes of all chunks when they fit GNNUDS_TAPE_MB
 * (16 GiB by default), otherwise only every chunk's input state, and the reverse sweep then re-runs one chunk at a
 * time with the tape on (checkpointing); it carries the state cotangents across the feedback
 * (h, q_us, q_ds and the edge outputs; the setting written into the edge state's last channel feeds grad_a; the flood
 * flag and the runoff carry no gradient).  x0 (B,seq_in,n_node,n_in), b (B,n_chunks*seq_out,n_node,b_in),
 * a (B,n_chunks*seq_out,n_act), ex0 (B,seq_in,n_edge,e_in) live where `where` says; the weight vectors / gamma
 * (n_chunks*seq_out) are HOST vectors as for gnnuds_objective. */
int gnnuds_grad_settings(gnnuds_handle* h, int32_t rule, const float* x0, const float* b, const float* a,
                         const float* ex0, int64_t B, int32_t n_chunks, const float* flood_w, const float* wwtp_w,
                         const float* rough_w, const float* gamma, float* obj, float* grad_a, int32_t where);

/* Event replay (main.py:306-382 test loop, dataloader.py:93-95 expand_seq): one rainfall event of T recorded steps -
 * states (T,N,4) = [h, cuminflow, outflow, lateral inflow], perfs (T,N) cumflooding (if_flood models, else NULL),
 * edge_states (T,M,4), settings (T,n_act) (act models, else NULL) - is uploaded ONCE and its T - seq_out teacher-forced
 * windows are built on the device by index arithmetic (the reference expands every series seq-fold on the host):
 * window w = states of steps w-seq_in..w-1 (zero-padded before the event start, q_us = cuminflow - lateral, flood flag =
 * cumflooding > 0), runoff of steps w+seq_out-seq.. (seq = max(seq_in,seq_out), first seq_out of them), settings of steps
 * w..w+seq_out-1.  gnnuds_replay runs Emulator.simulate over all windows: y (T-seq_out, seq_out, N, y_channels),
 * ey (T-seq_out, seq_out, M, ey_channels).  gnnuds_build_windows returns windows w0..w0+nw-1 themselves
 * (x (nw,seq_in,N,n_in), b (nw,seq_out,N,b_in), a (nw,seq_out,n_act) or NULL, ex (nw,seq_in,M,4)). */
int gnnuds_replay(gnnuds_handle* h, const float* states, const float* perfs, const float* edge_states, const float* settings,
                  int64_t T, float* y, float* ey, int32_t where);
int gnnuds_build_windows(gnnuds_handle* h, const float* states, const float* perfs, const float* edge_states,
                         const float* settings, int64_t T, int64_t w0, int64_t nw, float* x, float* b, float* a, float* ex,
                         int32_t where);

/* n_events recorded events in ONE call (the reference's test loop, main.py:306-382, walks them one by one): the windows
 * of all events form one sequence that is cut into workspace chunks regardless of event boundaries, a chunk's windows are
 * built on the device and run as one batch, and the predictions of chunk c leave through the pinned ring while chunk c+1
 * computes.  Host pointers only: states[i] (T[i],N,4), perfs[i] (T[i],N) (if_flood models; the array may be NULL
 * otherwise), edge_states[i] (T[i],M,4), settings[i] (T[i],n_act) (act models; else the array may be NULL); y[i] / ey[i]
 * as gnnuds_replay returns them for event i.  Bitwise the results of gnnuds_replay event by event. */
int gnnuds_replay_batch(gnnuds_handle* h, int32_t n_events, const float* const* states, const float* const* perfs,
                        const float* const* edge_states, const float* const* settings, const int64_t* T,
                        float* const* y, float* const* ey);

/* Pinned host buffers for callers that want truly asynchronous copies. */
int gnnuds_alloc_pinned(void** p, int64_t bytes);
int gnnuds_free_pinned(void* p);
/* Device buffers on the handle's device (for where = GNNUDS_MEM_DEVICE callers without torch). */
int gnnuds_alloc_device(gnnuds_handle* h, void** p, int64_t bytes);
int gnnuds_free_device(gnnuds_handle* h, void* p);
int gnnuds_memcpy(gnnuds_handle* h, void* dst, const void* src, int64_t bytes, int32_t kind /*1 H2D,2 D2H*/);
int gnnuds_synchronize(gnnuds_handle* h);

/* Kernel launches issued by this handle since creation (bench.py "gpu_launches"). */
int64_t gnnuds_launch_count(const gnnuds_handle* h);
/* Device time (ms, CUDA events on the handle's stream) of the last predict/rollout call's
   compute (copies excluded) and of its dominant GEMM kernels. */
int gnnuds_last_timing(const gnnuds_handle* h, float* total_ms, float* gemm_ms);
int gnnuds_set_profiling(gnnuds_handle* h, int32_t on);
/* With profiling level 2: summed device time and launch count of one kernel family in the last call.
   No reference counterpart (the reference only prints wall-clock "Emulation time", main.py:362). */
#define GNNUDS_FAM_GEMM 0     /* tc_rows_kernel / rows_gemm_f32: Dense, conv linear, Conv1D, GRU/LSTM projections */
#define GNNUDS_FAM_SPATIAL 1  /* sp_agg_kernel / sp_aggt_kernel / sp_conv_kernel: fused NodeEdge + conv + aggregation */
#define GNNUDS_FAM_GATHER 2   /* agg_large_kernel: CSR neighbour gather of the large-graph path */
#define GNNUDS_FAM_RECURRENT 3 /* gru_kernel / lstm_kernel: in-kernel time loop */
int gnnuds_kernel_timing(const gnnuds_handle* h, int32_t family, float* ms, int64_t* launches);

/* Test hook: stop the forward after `stage` (0 = off) and fetch the live node/edge
   streams of the first chunk.  which: 0 node stream, 1 edge stream. */
int gnnuds_debug_stop_after(gnnuds_handle* h, int32_t stage);
int gnnuds_debug_fetch(gnnuds_handle* h, int32_t which, float* out, int64_t max_floats,
                       int64_t* rows, int32_t* width);

/* Test hook for the tcgen05 building blocks (csrc/umma.cuh), one CTA, host buffers:
     mode 0: out(128,Rm)  = bf16(act[shift:shift+128, :Ca]) @ bf16(mat(Rm,Km=Ca))^T
             activations as the K-major A operand (with a row shift, as the causal Conv1D taps
             use it), the weight image as the K-major B operand;
     mode 1: out(128,Ca)  = bf16(mat(128,Km)) @ bf16(act[:Km, :Ca])
             the weight image as A, activations as the MN-major B operand (NodeEdge / aggregation).
   swap != 0 exchanges the leading/stride byte offsets of every descriptor (diagnostic). */
int gnnuds_debug_umma_selftest(gnnuds_handle* h, int32_t mode, const float* act, int32_t Ra, int32_t Ca,
                               const float* mat, int32_t Rm, int32_t Km, int32_t shift, int32_t swap,
                               float* out);

#ifdef __cplusplus
}
#endif
#endif /* GNNUDS_H */
