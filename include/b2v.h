/*
 * b2v.h — C-ABI of libb2v.so, the B200 (sm_100a) implementation of one hot path of
 * noorgme/Decentralised-Verification-and-Distributed-Execution-of-Large-Scale-Video-Diffusion-Models:
 * the chunked-coherent Zeroscope (UNet3D) denoising loop of Distribution/strategies and the
 * Proof-of-Inference Merkle commit / verify of InferNet/template/validator/proof.py.
 *
 * The reference has no FFI (it is pure Python on torch/diffusers/hashlib); each entry point below cites
 * the reference call site (file:line, relative to the reference root) whose work it replaces.
 * INTEGRATION.md shows the ctypes binding a reference maintainer would add.
 *
 * Conventions
 *   - every function returns 0 on success or a negative B2V_ERR_* code; b2v_last_error() gives the
 *     thread-local message. Nothing here falls back to the CPU: without an sm_100 device the calls fail.
 *   - pointers named *_dev are device pointers on the current CUDA device, *_host are host pointers.
 *   - `stream` is a cudaStream_t passed as void* (NULL = default stream). Launches are asynchronous.
 *   - no hidden device allocation in the per-step calls: the caller passes workspaces.
 *   - tensors use the reference's layouts at the boundary: latents are contiguous NCFHW [B,4,F,h,w]
 *     (chunk_only.py:110-113), text embeddings [B,77,1024] (chunk_only.py:62), leaf bytes are the
 *     C-order little-endian fp16 image of the latent (neurons/miner.py:203-204).
 */
#ifndef B2V_H_
#define B2V_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2V_VERSION 1

/* status codes */
#define B2V_OK 0
#define B2V_ERR_ARG (-1)     /* bad argument / unsupported shape */
#define B2V_ERR_CUDA (-2)    /* CUDA runtime / driver error */
#define B2V_ERR_DEVICE (-3)  /* no sm_100 device: there is no CPU fallback */
#define B2V_ERR_STATE (-4)   /* handle not finalised, missing weight ... */
#define B2V_ERR_WORKSPACE (-5)

/* dtype tags */
#define B2V_F32 0
#define B2V_F16 1
#define B2V_BF16 2

int b2v_version(void);
const char* b2v_last_error(void);
/* 0 when the current device is compute capability 10.x, B2V_ERR_DEVICE otherwise. */
int b2v_device_check(void);
/* number of kernels launched by this library in this process so far (bench.py: gpu_launches). */
uint64_t b2v_launch_count(void);

/* ------------------------------------------------------------------------------------------------
 * Proof-of-Inference commitment  (InferNet/neurons/miner.py:180-264, template/validator/proof.py:250-304)
 * ------------------------------------------------------------------------------------------------ */

/* Leaf hashing. digest[i] = SHA256( prefix_i ‖ A_i ‖ B_i ) where prefix_i is the 2-byte big-endian
 * timestep (miner.py:207 `t.to_bytes(2,'big') + z_bytes + eps_bytes`) when prefix_be16_host != NULL, or
 * empty (proof.py:256 `hashlib.sha256(leaf)` on a raw leaf). A_i / B_i are device buffers (B may be
 * NULL / 0 bytes). One independent Merkle–Damgård chain per leaf; mode 0 = auto, 1 = one CTA per chain
 * (few long leaves), 2 = one lane per chain (thousands of leaves), 3 = one warp per chain without shared memory
 * (a leaf hashed on a side stream BESIDE the UNet's resident one-CTA-per-SM kernels: the in-loop commitment).
 * a_ptrs_host/b_ptrs_host: host arrays of n device pointers; lens in bytes; digests_dev: [n][32]. */
int b2v_sha256_leaves(const uint16_t* prefix_be16_host, const void* const* a_ptrs_host,
                      const uint64_t* a_lens_host, const void* const* b_ptrs_host,
                      const uint64_t* b_lens_host, int n, uint8_t* digests_dev, int mode, void* stream);

/* Tree levels + sibling paths (miner.py:216-249 == proof.py:260-286): parent = SHA256(min(l,r)‖max(l,r))
 * with byte-wise comparison, an odd tail node is paired with itself, the sibling of a missing node is
 * the node itself. depth = b2v_merkle_depth(n) = number of levels below the root (0 for n==1).
 * leaf_hashes_dev [n][32] -> root_dev [32], paths_dev [n][depth][32] (may be NULL).
 * scratch_dev: at least b2v_merkle_scratch_bytes(n) bytes. */
int b2v_merkle_depth(int n);
size_t b2v_merkle_scratch_bytes(int n);
int b2v_merkle_build(const uint8_t* leaf_hashes_dev, int n, uint8_t* root_dev, uint8_t* paths_dev,
                     void* scratch_dev, void* stream);

/* b2v_merkle_build without a scratch argument and without any device allocation. It has no counterpart in the
 * reference (which builds its tree in Python lists on the host); it exists for hosts that cannot hand this library a
 * scratch buffer because they are not allowed to allocate one at the point of the call:
 *   - a miner that commits its trajectory from a helper thread while the denoising loop is running on the same GPU:
 *     cudaMalloc / cudaFree stall until the device is idle, so an allocation there stops the loop it runs beside;
 *   - a serving process that has given all of HBM to its own pool (UNet workspace, VAE, leaf store) and reports
 *     any allocation made behind its back as a leak, or simply has nothing left to give.
 * The tree levels are kept in an array that is part of libb2v.so's device image: about 257 KiB on every GPU the
 * library is loaded on, held until the process exits. Inputs and outputs are the caller's, as in b2v_merkle_build
 * (leaf_hashes_dev [n][32] -> root_dev [32], paths_dev [n][depth][32] or NULL), and the results are the same bytes.
 *   - n must be 1 .. B2V_MERKLE_STATIC_MAX_LEAVES (also returned by b2v_merkle_static_max_leaves()); anything larger
 *     is refused with B2V_ERR_ARG and a message naming the limit, and needs b2v_merkle_build.
 *   - the first call on each device looks up the array's device address (once per device, with that device
 *     current) and remembers it; make that first call outside any stream capture, e.g. at start-up on a dummy leaf.
 *   - all calls on a device share the one array. Nothing in it has to start at zero: each level is written in full
 *     before it is read, so the previous call's contents are overwritten, never consumed. Two calls whose work
 *     overlaps on the GPU (different streams, or different host threads on different streams) race on the array and
 *     give wrong roots: issue them on one stream, or order them with events.
 *   - it calls no cudaMalloc, cudaFree, stream-ordered allocator or other allocating routine. */
#define B2V_MERKLE_STATIC_MAX_LEAVES 4096
int b2v_merkle_static_max_leaves(void);
int b2v_merkle_build_static(const uint8_t* leaf_hashes_dev, int n, uint8_t* root_dev, uint8_t* paths_dev,
                            void* stream);

/* Batched leaf verification (proof.py:291-304 verify_merkle_leaf): for each of m leaves fold
 * cur = SHA256(min(cur,sib)‖max(cur,sib)) over `depth` siblings and compare with root.
 * ok_dev[i] = 1/0. */
int b2v_merkle_verify(const uint8_t* leaf_hashes_dev, const uint8_t* paths_dev, int m, int depth,
                      const uint8_t* root_dev, uint8_t* ok_dev, void* stream);

/* The same fold for leaves of DIFFERENT trees (a validator round over many miners, neurons/validator.py:768-790):
 * leaf i is checked against roots_dev[i] ([m][32]). */
int b2v_merkle_verify_roots(const uint8_t* leaf_hashes_dev, const uint8_t* paths_dev, int m, int depth,
                            const uint8_t* roots_dev, uint8_t* ok_dev, void* stream);

/* Spot-check DDIM re-check on device (proof.py:307-392 run_unet_step and proof.py:419-626
 * verify_temporal_coherence_spot_check, default-config DDIMScheduler branch, Appendix C of SURVEY.md).
 * For each of m leaf pairs: z_i/eps_i/z_j are fp16 [1,C,F,h,w] device buffers of which only frame 0 is
 * examined (proof.py:330-338); prev = sqrt(a_prev)*clamp((z-sqrt(1-a_t)*eps)/sqrt(a_t),-1,1)
 * + sqrt(1-a_prev)*eps evaluated with the reference's per-op fp16 rounding.
 * flags_dev[i] bit0 = all finite, bit1 = max|prev| < 10 (run_unet_step verdict = bit0&bit1),
 * bit2 = allclose(prev, z_j, rtol=1, atol=1) (temporal-coherence verdict; z_j may be NULL -> bit2=1). */
int b2v_spotcheck_ddim(const void* const* z_ptrs_host, const void* const* eps_ptrs_host,
                       const void* const* znext_ptrs_host, const float* a_t_host,
                       const float* a_prev_host, int m, int C, int F, int64_t hw, uint8_t* flags_dev,
                       void* stream);

/* ------------------------------------------------------------------------------------------------
 * Denoising step glue  (Distribution/strategies/chunk_only.py:65-74, fsdp_chunked_coherent.py:129-143)
 * ------------------------------------------------------------------------------------------------ */

/* Fused classifier-free-guidance combine + DDIM(eta=0) update + next UNet input.
 *   eps   = cfg ? u + gs*(c-u) : eps2[0]          (chunk_only.py:72-73; miner.py:581-586 has no CFG)
 *   x0    = (lat - sqrt(1-a_t)*eps)/sqrt(a_t) ; clip to [-1,1] when clip_sample
 *   lat'  = sqrt(a_prev)*x0 + sqrt(1-a_prev)*eps  (diffusers DDIMScheduler.step, eta=0)
 *   x'    = [lat' (+ ctx_w*ctx broadcast over frames)] duplicated n_dup times   (chunk_only.py:68,
 *           fsdp_chunked_coherent.py:134-137) — the next step's UNet input, may be NULL
 *   leaf_z_f16 / leaf_eps_f16 (may be NULL): fp16 images of lat (the step's input) and eps, i.e. the
 *           bytes miner.py:203-204 hashes.
 * lat/lat_out: [1,C,F,h,w]; eps2: [2 or 1,C,F,h,w]; ctx: [1,C,1,h,w] or NULL; all of dtype `dtype`.
 * Math is fp32 with one rounding at the store. */
int b2v_cfg_ddim_step(const void* lat_dev, const void* eps2_dev, const void* ctx_dev, float ctx_w,
                      int cfg, float guidance_scale, float a_t, float a_prev, int clip_sample,
                      void* lat_out_dev, void* unet_in_next_dev, int n_dup, void* leaf_z_f16_dev,
                      void* leaf_eps_f16_dev, int dtype, int C, int F, int64_t hw, void* stream);

/* First UNet input of a chunk: x = dup(lat (+ ctx_w*ctx)) (fsdp_chunked_coherent.py:133-137). */
int b2v_make_unet_input(const void* lat_dev, const void* ctx_dev, float ctx_w, void* unet_in_dev,
                        int n_dup, int dtype, int C, int F, int64_t hw, void* stream);

/* Global context = mean over frames of the base noise (fsdp_chunked_coherent.py:112-118).
 * base [1,C,T,h,w] -> ctx [1,C,1,h,w], fp32 accumulate, one rounding. */
int b2v_context_mean(const void* base_dev, void* ctx_dev, int dtype, int C, int T, int64_t hw, void* stream);

/* Overlap blend (chunk_only.py:136-150), reproduced exactly: `full` is an accumulator of the latents'
 * own dtype that is rounded after every chunk's `+= latc*w` in gather order, weight is fp32, the result
 * is full/clamp(weight,1e-6) in fp32. chunk_ptrs_host[i] -> device (or peer-mapped) [1,C,len_i,h,w];
 * w_host: concatenated per-chunk per-frame fp32 weights (sum len_i entries, built by the host from
 * torch.linspace like the reference). Outputs: lat_out_f32_dev [1,C,T,h,w] fp32 and, optionally,
 * full_out_dev (dtype) for inspection. */
int b2v_blend_overlap(const void* const* chunk_ptrs_host, const int* starts_host, const int* ends_host,
                      int n_chunks, const float* w_host, int C, int T, int64_t hw, int dtype,
                      float* lat_out_f32_dev, void* full_out_dev, void* stream);

/* Wire format of the commitment (InferNet/neurons/miner.py:643-657 and the spot-check reveal 417-423:
 * `base64.b64encode(z_bytes).decode('ascii')` for every leaf image). Encodes n_rows rows of row_bytes bytes
 * (row i at src_dev + i*row_stride) into dst_dev, row i at dst_dev + i*b2v_base64_len(row_bytes): RFC 4648
 * standard alphabet with '=' padding, bit-identical to Python's base64.b64encode. */
uint64_t b2v_base64_len(uint64_t nbytes);
int b2v_base64_rows(const void* src_dev, int64_t n_rows, uint64_t row_bytes, uint64_t row_stride,
                    void* dst_dev, void* stream);
/* The validator's side (neurons/validator.py:783-786 `base64.b64decode(z_b)` per revealed leaf): n_rows rows of
 * row_chars characters (a multiple of 4; row i at src_dev + i*row_stride) -> bytes at dst_dev + i*dst_stride
 * (3*row_chars/4 bytes written, padded positions as zero). status_dev[i] = number of '=' padding characters (0..2),
 * | 0x100 when the row is not canonical base64 (character outside the alphabet, '=' before the end): Python's
 * permissive decoder skips such characters, so the caller decodes those rows on the host to keep its semantics. */
int b2v_base64_decode_rows(const void* src_dev, int64_t n_rows, uint64_t row_chars, uint64_t row_stride,
                           void* dst_dev, uint64_t dst_stride, int* status_dev, void* stream);

/* ------------------------------------------------------------------------------------------------
 * UNet3DConditionModel forward  ([UPSTREAM diffusers] pipe.unet(x, t, encoder_hidden_states=emb).sample,
 * chunk_only.py:71, fsdp_chunked_coherent.py:140, miner.py:581)
 * ------------------------------------------------------------------------------------------------ */

typedef struct b2v_unet3d_config {
  int in_channels;            /* 4 */
  int out_channels;           /* 4 */
  int block_out_channels[4];  /* 320,640,1280,1280 */
  int layers_per_block;       /* 2 */
  int cross_attention_dim;    /* 1024 */
  int attention_head_dim;     /* 64 (heads = C/64) */
  int norm_num_groups;        /* 32 */
  float norm_eps;             /* 1e-5 */
  int transformer_in_heads;   /* 8 */
  int text_len;               /* 77 */
  int act_dtype;              /* element type of activations and GEMM operands inside the network: B2V_BF16 (also when 0)
                                 or B2V_F16 — the reference's own arithmetic (chunk_only.py:44-47 torch_dtype=float16,
                                 miner.py:83-86). Same tensor-core rate; fp16 carries 3 more mantissa bits. */
} b2v_unet3d_config;

typedef struct b2v_unet3d b2v_unet3d;

int b2v_unet3d_create(const b2v_unet3d_config* cfg, b2v_unet3d** out);
void b2v_unet3d_destroy(b2v_unet3d* h);
/* Feed one tensor of the diffusers state_dict (name as in the checkpoint, e.g.
 * "down_blocks.0.resnets.0.conv1.weight"); data may be host or device memory of dtype `dtype`.
 * The library repacks it for its kernels (bf16 GEMM operands, fp32 biases / norm affine). */
int b2v_unet3d_load_tensor(b2v_unet3d* h, const char* name, const void* data, int dtype, int ndim,
                           const int64_t* shape);
/* Verify that every parameter was supplied and freeze the weight set. */
int b2v_unet3d_finalize(b2v_unet3d* h);
/* number of parameters loaded (sanity anchor: 1 411 2xx xxx for Zeroscope). */
int64_t b2v_unet3d_param_count(const b2v_unet3d* h);
size_t b2v_unet3d_workspace_bytes(const b2v_unet3d* h, int B, int F, int H, int W);
/* sample/out: [B,Cin,F,H,W] NCFHW of `io_dtype`; ehs: [B,text_len,cross_attention_dim] of `io_dtype`
 * (replicated over frames internally, never materialised). */
int b2v_unet3d_forward(b2v_unet3d* h, const void* sample_dev, float timestep, const void* ehs_dev,
                       void* out_dev, int io_dtype, int B, int F, int H, int W, void* workspace_dev,
                       size_t workspace_bytes, void* stream);
/* algorithmic FLOPs of one forward (GEMM/conv/attention only, MAC = 2), the roofline numerator. */
double b2v_unet3d_flops(const b2v_unet3d* h, int B, int F, int H, int W);

/* ------------------------------------------------------------------------------------------------
 * Building-block ops (used by the UNet; exported so tests/ can check each kernel against the oracle).
 * Activations are channels-last 2-byte rows: [rows, C] with rows = (b, f, y, x) flattened. `act_dtype` (B2V_BF16 or
 * B2V_F16) names the element type of every activation / weight operand of the call; accumulation is always fp32.
 * ------------------------------------------------------------------------------------------------ */

/* epilogue flags for b2v_op_gemm */
#define B2V_EPI_GEGLU 1      /* W rows are [value | gate] interleaved per 128-column tile pair; out has N/2 columns */
#define B2V_EPI_GELU 2

/* out[M,N] = A[M,K] · W[N,K]^T (+bias[N]) (+rowbias[row/rows_per_group, N]) (+residual[M,N]); act_dtype in/out,
 * fp32 accumulate on tcgen05. taps/tap_row_stride: 1 tap = plain GEMM; 3 taps = temporal (3,1,1) conv
 * where tap t reads row r+(t-1)*tap_row_stride inside a group of rows_per_batch rows, zero outside
 * (K = taps*Kc, W laid out [N][tap][Kc]). */
int b2v_op_gemm(const void* a_dev, const void* w_dev, const float* bias_dev, const float* rowbias_dev,
                int64_t rows_per_group, const void* residual_dev, void* out_dev, int64_t M, int N, int Kc,
                int taps, int64_t tap_row_stride, int64_t rows_per_batch, int flags, int act_dtype, void* stream);

/* 3x3 convolution, padding 1, stride 1 or 2, NHWC, weights [Cout][3][3][Cin], implicit GEMM
 * (TMA im2col -> tcgen05). mode 0 = TMA im2col, 1 = explicit im2col matrix in `scratch_dev` then GEMM. */
size_t b2v_op_conv3x3_scratch_bytes(int N, int H, int W, int Cin, int stride, int mode);
int b2v_op_conv3x3(const void* x_dev, const void* w_dev, const float* bias_dev, const float* rowbias_dev,
                   int64_t rows_per_group, const void* residual_dev, void* out_dev, int N, int H, int W,
                   int Cin, int Cout, int stride, int mode, void* scratch_dev, int act_dtype, void* stream);

/* GroupNorm (+SiLU) over channels-last rows: stats per (batch item, group) over rows_per_item rows x
 * C/groups channels, then y = (x-mean)*rstd*gamma+beta (silu optional). x may be the channel concat of
 * two tensors (x2_dev != NULL, C = C1 + C2). scratch_dev: >= b2v_op_group_norm_scratch_bytes() bytes (per-CTA
 * partial sums; the reduction uses no atomics, so results are bit-reproducible run to run). */
size_t b2v_op_group_norm_scratch_bytes(int64_t n_items, int64_t rows_per_item, int groups);
int b2v_op_group_norm(const void* x_dev, int C1, const void* x2_dev, int C2, const float* gamma_dev,
                      const float* beta_dev, void* y_dev, int64_t n_items, int64_t rows_per_item,
                      int groups, float eps, int silu, void* scratch_dev, int act_dtype, void* stream);
int b2v_op_layer_norm(const void* x_dev, const float* gamma_dev, const float* beta_dev, void* y_dev,
                      int64_t rows, int C, float eps, int act_dtype, void* stream);

/* GroupNorm statistics out of the PRODUCING GEMM / conv (ResnetBlock2D conv1 -> norm2, TemporalConvLayer conv_i ->
 * norm_{i+1}, proj_out -> the next block's norm: the statistics pass over the activation disappears).
 * b2v_op_gemm_gnstats / b2v_op_conv3x3_gnstats are b2v_op_gemm / b2v_op_conv3x3 (stride 1, TMA im2col) whose epilogue
 * also writes gn_stats_dev[M/32][gn_groups][2] = fp32 {sum, sum of squares} of the output per 32-row block and group
 * (fixed-order, bit-reproducible). Shapes qualify when b2v_gn_stats_ok() is 1: the GEMM's tile width (160 for the UNet's
 * 320 / 640 / 1280 channels, 256 or 128 for the VAE decoder's 512 / 256 / 128) holds whole groups per half tile (10 / 20 /
 * 40 channels per group at 160; 16 / 32 at 256; 4 at 128) and M is whole 256-row units; otherwise the calls fail with
 * B2V_ERR_ARG. The consumer folds the partials per item in 32-row blocks, so rows_per_item must be a multiple of 32.
 * b2v_op_group_norm_pre is b2v_op_group_norm (no concat) fed with such statistics instead of its own first pass. */
int b2v_gn_stats_ok(int64_t M, int N, int64_t rows_per_batch, int groups);
int b2v_op_gemm_gnstats(const void* a_dev, const void* w_dev, const float* bias_dev, const float* rowbias_dev,
                        int64_t rows_per_group, const void* residual_dev, void* out_dev, int64_t M, int N, int Kc,
                        int taps, int64_t tap_row_stride, int64_t rows_per_batch, int act_dtype, float* gn_stats_dev,
                        int gn_groups, void* stream);
int b2v_op_conv3x3_gnstats(const void* x_dev, const void* w_dev, const float* bias_dev, const float* rowbias_dev,
                           int64_t rows_per_group, const void* residual_dev, void* out_dev, int N, int H, int W,
                           int Cin, int Cout, int act_dtype, float* gn_stats_dev, int gn_groups, void* stream);
/* LayerNorm fused across two linear layers (BasicTransformerBlock: norm1 -> to_q/k/v, norm2 -> to_q, norm3 -> GEGLU):
 * the GEMM that PRODUCES the hidden states (proj_in, to_out + residual) also writes, per output row, the {sum, sum of
 * squares} of the row over each epilogue warp's columns (b2v_op_gemm_rowstats: row_stats_dev[M][b2v_row_stats_slots()]
 * float2); the GEMM that CONSUMES them reads the raw hidden states with W' = W diag(gamma) and finishes the
 * normalisation in its epilogue: out = rstd (acc - mean csum) + bias', csum[n] = sum_k W'[n][k] (of the rounded W'),
 * bias' = b + W beta (b2v_op_gemm_lnin; flags may carry B2V_EPI_GEGLU with W', bias', csum packed like W).
 * The LayerNorm pass (one read + one write of the activation) disappears. */
int b2v_row_stats_slots(int64_t M, int N, int64_t rows_per_batch);
int b2v_op_gemm_rowstats(const void* a_dev, const void* w_dev, const float* bias_dev, const void* residual_dev,
                         void* out_dev, int64_t M, int N, int Kc, int act_dtype, float* row_stats_dev, void* stream);
int b2v_op_gemm_lnin(const void* a_raw_dev, const void* w_gamma_dev, const float* bias_folded_dev, const float* csum_dev,
                     const float* ln_stats_dev, int ln_slots, float eps, void* out_dev, int64_t M, int N, int Kc,
                     int flags, int act_dtype, void* stream);
int b2v_op_group_norm_pre(const void* x_dev, int C, const float* gamma_dev, const float* beta_dev, void* y_dev,
                          int64_t n_items, int64_t rows_per_item, int groups, float eps, int silu, void* scratch_dev,
                          const float* pre_dev, int act_dtype, void* stream);

/* softmax(Q K^T * scale) V, head_dim 64. q: [nb, Sq, ldq] rows, head h at column h*64;
 * k/v: [nb / kv_group, Skv, ldk]; out [nb, Sq, ldo]. (tcgen05 flash attention.) */
int b2v_op_attention(const void* q_dev, int64_t ldq, const void* k_dev, const void* v_dev, int64_t ldkv,
                     void* out_dev, int64_t ldo, int nb, int heads, int Sq, int Skv, int kv_group,
                     float scale, int act_dtype, void* stream);
/* attention over the frame axis: tokens [B, F, HW, *], sequence = F (<= 32), head_dim 64. */
int b2v_op_temporal_attention(const void* qkv_dev, int64_t ld, int q_off, int k_off, int v_off,
                              void* out_dev, int64_t ldo, int B, int F, int64_t HW, int heads,
                              float scale, int act_dtype, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Frame decode after the loop  ([UPSTREAM diffusers] pipe.vae.decode(z / 0.18215).sample and the uint8 frame
 * conversion, chunk_only.py:152-158, fsdp_chunked_coherent.py:219-227, miner.py:594-600). The decoder's
 * convolutions / projections / GroupNorms are the ops above; these are the remaining pieces. The host-side
 * composition is b2v/vae.py (AutoencoderKL.decode / decode_frames).
 * ------------------------------------------------------------------------------------------------ */

/* x[r, 0:cols] = softmax(scale * x[r, 0:cols]) in place, bf16 rows of leading dimension ld, fp32 arithmetic:
 * the mid-block attention of the decoder is ONE head of width 512 (S = Q K^T and O = P V are b2v_op_gemm calls). */
int b2v_op_softmax_rows(void* x_dev, int64_t rows, int cols, int64_t ld, float scale, void* stream);
/* Decoder output rows [n_px, ld] bf16 (channels 0..2 = RGB in [-1,1]) -> out [n_px, 3] bytes:
 * (v*0.5+0.5).clamp(0,1)*255 truncated (chunk_only.py:157-158 `.byte()`); frames come out as [H,W,3]. */
int b2v_op_frames_u8(const void* rows_dev, int64_t ld, int64_t n_px, uint8_t* out_dev, void* stream);
/* Nearest-neighbour x2 (Upsample2D, F.interpolate(scale_factor=2, mode="nearest")): [N,H,W,C] -> [N,2H,2W,C] bf16. */
int b2v_op_upsample2x(const void* x_dev, void* y_dev, int N, int H, int W, int C, void* stream);
/* Latents [B,C,F,h,w] of `io_dtype` -> channels-last rows [(b,f,y,x), Cpad] of act_dtype, channels >= C zero: the
 * layout the tensor-core conv_in / post_quant_conv read. */
int b2v_op_latent_rows(const void* x_dev, int io_dtype, void* rows_dev, int B, int C, int F, int64_t hw, int Cpad,
                       int act_dtype, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B2V_H_ */
