/*
 * pcc_b200.h — C ABI of the B200-native learned point-cloud codec hot path.
 *
 * Every entry point is `extern "C"`, takes plain device pointers + element
 * counts + a CUDA stream, never allocates, and returns 0 on success or a
 * negative PCC_E_* code (message via pcc_last_error()).  Buffers whose size is
 * data dependent are sized by the caller to a host-visible upper bound; the
 * true count comes back through a device counter (int32_t*).
 *
 * Each function cites the reference call site (file:line under the reference
 * tree) it replaces.  The reference reaches that arithmetic through
 * MinkowskiEngine / CompressAI Python bindings; INTEGRATION.md shows the
 * ctypes stub a maintainer would add on the reference side.
 *
 * Coordinate representation used across the ABI: a coordinate (b,x,y,z)
 * (b >= 0, x,y,z in int16 range) is one uint64 "key":
 *     key = b << 48 | morton3(x + 32768, y + 32768, z + 32768)
 * where bit i of x lands on key bit 3i, y on 3i+1, z on 3i+2.  Coordinate sets
 * are always stored sorted ascending by key and duplicate free; row r of a
 * feature matrix belongs to keys[r].
 */
#ifndef PCC_B200_H
#define PCC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* pcc_stream_t; /* cudaStream_t */

#define PCC_OK 0
#define PCC_E_INVALID (-1)   /* bad argument */
#define PCC_E_CUDA (-2)      /* CUDA runtime error (launch / copy) */
#define PCC_E_WORKSPACE (-3) /* workspace too small */
#define PCC_E_UNSUPPORTED (-4)

int pcc_version(void);
const char* pcc_last_error(void);
/* Number of kernels this library has launched in this process (bench.py's gpu_launches). */
int64_t pcc_launch_count(void);

/* Device fault word.  Kernels never trap (a trap is a sticky error that ends the CUDA context of the whole
 * encoder / decoder service, sender/encoder/encoder.py:50-75, receiver/decoder/decoder.py:41-62 keep one process
 * alive for the session): a kernel that cannot go on records a PCC_FAULT_* code in a per-device word and returns
 * early, leaving its outputs unwritten.  pcc_fault_fetch copies the word of the current device into dst_dev (device
 * int32) and clears it, stream ordered; pcc_pack_strings also delivers and clears it in its header.  A non-zero
 * value means the results of the calls since the last fetch must be discarded. */
#define PCC_FAULT_BLOCKMAP_FULL 1
#define PCC_FAULT_PIPELINE_TIMEOUT 2
int pcc_fault_fetch(int32_t* dst_dev, pcc_stream_t s);
/* Every deferred check of a call in one launch: dst_dev[i] = *srcs[i] (srcs: HOST array of n < 48 device addresses of
 * int32 counters / status words), dst_dev[n] = the fault word, fetched and cleared, when with_fault != 0.  The caller
 * then reads dst_dev with one copy (decompress()'s single synchronisation). */
int pcc_fetch_words(const int32_t* const* srcs_dev, int32_t n, int32_t with_fault, int32_t* dst_dev, pcc_stream_t s);
const char* pcc_fault_string(int32_t code);
/* Test hook: what = 1 makes the next k = 3 pcc_spconv_umma launch of this process drop one pipeline hand-off (and spin
 * briefly instead of for seconds), so that the timeout path can be exercised; 0 disarms. */
int pcc_debug_inject(int32_t what);
/* Debug hook: the next cout = 32 pcc_spconv_octet launch records per-stage timestamps (globaltimer, ns) of its CTA 0 into
 * buf (device memory, pcc_debug_trace_words() int64 words): tools/octet_trace.py prints the pipeline timeline. */
int64_t pcc_debug_trace_words(void);
int pcc_debug_trace(long long* buf);

/* ------------------------------------------------------------------------------------------
 * (1) coordinates, hash, kernel maps
 * ------------------------------------------------------------------------------------------ */

/* ME.SparseTensor(coordinates=float/int [N,4]) coordinate conversion: floor + pack.
 * Replaces the coordinate quantisation of ME.SparseTensor at codec_pipeline.py:262-266,
 * codec_parallel.py:296-301. */
int pcc_keys_from_coords_i32(const int32_t* coords, int64_t n, uint64_t* keys, pcc_stream_t s);
int pcc_keys_from_coords_f32(const float* coords, int64_t n, uint64_t* keys, pcc_stream_t s);
/* Raw capturer frame (int16 points [n,3], float64 colours [n,3], one batch index) -> keys and
 * (1,r,g,b) float features; fuses stack_tensors + the cat(ones, colors) of
 * codec_pipeline.py:257-260 (shared/utils.py:10-42). */
int pcc_ingest_frame(const int16_t* points, const double* colors, int64_t n, int32_t batch,
                     uint64_t* keys, float* feats4, pcc_stream_t s);
/* The same in two steps, so that the coordinate work (sort, pyramid, kernel maps) can start while the
 * colours -- 80 % of a raw frame's bytes -- are still on their way to the device: keys from the points
 * alone, then feature rows bf16 (1,r,g,b,0...) [m, c_dst] gathered through src_row (the first-wins rows
 * of pcc_sort_unique; NULL = identity; same rounding as pcc_ingest_frame + pcc_rows_f32_to_bf16). */
int pcc_ingest_points(const int16_t* points, int64_t n, int32_t batch, uint64_t* keys, pcc_stream_t s);
int pcc_ingest_features(const double* colors, const int32_t* src_row, int64_t m, void* feats_bf16,
                        int c_dst, pcc_stream_t s);
/* The same from colours already rounded to bf16 [n,3] (PCC_STAGE_F64_TO_BF16 staging). */
int pcc_ingest_features_bf16(const void* colors_bf16, const int32_t* src_row, int64_t m, void* feats_bf16,
                             int c_dst, pcc_stream_t s);
/* Decoder output packing (pack_batches, codec_parallel.py:474-502): keys -> int32 [n,3] (x,y,z) and
 * colours [n,3] with NaN -> 0 and clip(c * 255, 0, 255) / 255 evaluated in fp32.  Either output may
 * be NULL (the decoder sends the points to the host while the colour head still runs). */
int pcc_unpack_output(const uint64_t* keys, const float* feats3, int64_t n, int32_t* points, float* colors,
                      pcc_stream_t s);
/* The client's playout buffer straight from the decoder output (receiver/client/client.py:139-146, SURVEY
 * 8f row f-4): points + offset as float32 [n,3] and colours as uint8 [n,3] = trunc(255 * colour), colour as
 * pcc_unpack_output leaves it -- 15 bytes per point to bring to the host instead of 24. */
int pcc_pack_visualizer(const uint64_t* keys, const float* feats3, int64_t n, float offset, float* points_f32,
                        uint8_t* colors_u8, pcc_stream_t s);
/* keys -> int32 [n,4] (b,x,y,z): the `.C` attribute. */
int pcc_coords_from_keys(const uint64_t* keys, int64_t n, int32_t* coords, pcc_stream_t s);

/* Sort + de-duplicate (first occurrence wins): the hash insert of ME.SparseTensor construction.
 * src_row[i] (optional) = input row that output row i was taken from. */
int64_t pcc_sort_unique_workspace_bytes(int64_t n);
int pcc_sort_unique(const uint64_t* keys_in, int64_t n, uint64_t* keys_out, int32_t* src_row,
                    int32_t* n_out_dev, void* ws, int64_t ws_bytes, pcc_stream_t s);
/* Same, for keys whose bits [key_bits, 64) are zero (frame index < 2^(key_bits - 48)): fewer radix passes. */
int pcc_sort_unique_bits(const uint64_t* keys_in, int64_t n, int key_bits, uint64_t* keys_out, int32_t* src_row,
                         int32_t* n_out_dev, void* ws, int64_t ws_bytes, pcc_stream_t s);

/* Stride-2 convolution output coordinates: c' = floor(c / t') * t' (MinkowskiConvolution with
 * stride 2 inside g_a / h_a / g_s.down_conv: codec_pipeline.py:273,287; codec_parallel.py:302-303).
 * parent_row[i] = output row of input row i.  log2_stride = log2(t'). */
int64_t pcc_coords_downsample_workspace_bytes(int64_t n);
int pcc_coords_downsample(const uint64_t* keys_in, int64_t n, int log2_stride, uint64_t* keys_out,
                          int32_t* parent_row, int32_t* n_out_dev, void* ws, int64_t ws_bytes,
                          pcc_stream_t s);
/* pcc_coords_downsample for callers that may not allocate device memory (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same inputs, outputs and results as pcc_coords_downsample, but the scratch (run flags +
 * scan storage) is a file-scope `static __device__` array inside the library (4 MiB + 256 KiB, held by every GPU
 * that loads the library until exit) instead of a caller-owned workspace; the call performs no cudaMalloc /
 * cudaFree and calls no allocating routine.  Limit: n <= pcc_coords_downsample_static_max_rows() (1 Mi rows);
 * a larger n returns PCC_E_INVALID with a message naming the limit.  All calls on one device share the array:
 * it keeps the previous call's contents (nothing in it has to start at zero - every word a call reads it has
 * written first), and two calls that overlap on the GPU (different streams, or different host threads on
 * different streams) RACE; serialise them on one stream. */
int64_t pcc_coords_downsample_static_max_rows(void);
int pcc_coords_downsample_static(const uint64_t* keys_in, int64_t n, int log2_stride, uint64_t* keys_out,
                                 int32_t* parent_row, int32_t* n_out_dev, pcc_stream_t s);
/* The whole stride pyramid of a coordinate set in one call (the output coordinates of the chained
 * stride-2 convolutions of g_a and h_a, codec_pipeline.py:273,287, and of g_s.down_conv x2,
 * codec_parallel.py:302-303): level l (0-based) holds the keys at stride 2^(log2_stride+l+1) in
 * keys_out[l*n ...], parent_row[l*n + i] = row of level l-1's (or the input's) row i in level l, and
 * counts_dev[l] its size.  Counts stay on the device between levels, so the caller syncs once for all
 * of them.  frame_offsets_dev: [(levels+1)][num_frames+1] per-frame row ranges of the input and of
 * every level (num_frames = 0: skipped). */
int64_t pcc_coords_pyramid_workspace_bytes(int64_t n);
int pcc_coords_pyramid(const uint64_t* keys_in, int64_t n, int log2_stride, int levels, uint64_t* keys_out,
                       int32_t* parent_row, int32_t* counts_dev, int32_t num_frames,
                       int32_t* frame_offsets_dev, void* ws, int64_t ws_bytes, pcc_stream_t s);
/* Generative transposed convolution coordinates (k=2, s=2): 8 children per input voxel; child row
 * 8*p + j, j = dx + 2dy + 4dz  (MinkowskiGenerativeConvolutionTranspose inside g_s / h_s:
 * codec_parallel.py:376,469).  log2_stride = log2 of the OUTPUT tensor stride. */
int pcc_coords_children(const uint64_t* keys_in, int64_t n, int log2_stride, uint64_t* keys_out,
                        pcc_stream_t s);

/* Open-addressing hash (64-bit keys, windows of 4 slots probed by 4 cooperating lanes).
 * capacity must be a power of two >= 2n. */
int pcc_hash_build(const uint64_t* keys, int64_t n, uint64_t* table_keys, int32_t* table_vals,
                   int64_t capacity, pcc_stream_t s);
/* features_at_coordinates (codec_pipeline.py:401, codec_parallel.py:387): row of each query key or -1. */
int pcc_hash_query(const uint64_t* queries, int64_t nq, const uint64_t* table_keys,
                   const int32_t* table_vals, int64_t capacity, int32_t* rows, pcc_stream_t s);
/* Kernel map of a k=3, stride-1 convolution on tensor stride 2^log2_stride:
 * nbr[j*n + o] = row of (o + d_j) or -1, j = (dx+1) + 3(dy+1) + 9(dz+1).  Voxel-hash version
 * (26 cooperative probes per row); kept as the independent comparator of pcc_kmap_k3_blocks. */
int pcc_kmap_k3(const uint64_t* keys, int64_t n, int log2_stride, const uint64_t* table_keys,
                const int32_t* table_vals, int64_t capacity, int32_t* nbr, pcc_stream_t s);
/* Block table: the same open-addressing hash keyed on 4x4x4-voxel Morton blocks
 * (key >> (3*log2_stride + 6)); value = {64-bit occupancy mask, first row}.  Rows of a block are
 * contiguous in the sorted key set, so row(voxel idx) = base + popc(mask & ((1 << idx) - 1)).
 * table_vals16: capacity x 16 bytes.  capacity: power of two >= 2 x the number of occupied blocks
 * (<= n; the coordinates two strides coarser bound it, which keeps the table small enough to stay
 * in L2); a table that turns out too small traps (CUDA error) instead of dropping blocks.
 * pcc_kmap_k3_blocks builds the same [27, n] kernel map as pcc_kmap_k3 with <= 8 block probes per
 * row (a 3x3x3 neighbourhood spans at most 2x2x2 blocks). */
int pcc_blockmap_build(const uint64_t* keys, int64_t n, int log2_stride, uint64_t* table_keys,
                       void* table_vals16, int64_t capacity, pcc_stream_t s);
int pcc_kmap_k3_blocks(const uint64_t* keys, int64_t n, int log2_stride, const uint64_t* table_keys,
                       const void* table_vals16, int64_t capacity, int32_t* nbr, int64_t nbr_ld,
                       pcc_stream_t s);
/* Kernel map of a k=2, stride-2 convolution: nbr[j*nbr_ld + p] = child row with octant j or -1
 * (row pitch nbr_ld >= n_parent; see pcc_spconv_umma for why the pitch is padded). */
/* Measurement helper: count_dev (device int64) = number of entries >= 0 in a kernel map [kvol][nbr_ld] restricted
 * to its first n_out columns -- the pair count P of a sparse convolution (flops = 2 P Cin Cout). */
int pcc_kmap_count_pairs(const int32_t* nbr, int64_t nbr_ld, int64_t n_out, int kvol, int64_t* count_dev, pcc_stream_t s);
int pcc_kmap_down2(const uint64_t* child_keys, const int32_t* parent_row, int64_t n_child,
                   int log2_child_stride, int32_t* nbr, int64_t n_parent, int64_t nbr_ld,
                   pcc_stream_t s);
/* Per-frame row ranges: offsets[b] = first row with batch index >= b, b = 0..num_frames. */
int pcc_frame_offsets(const uint64_t* keys, int64_t n, int32_t num_frames, int32_t* offsets,
                      pcc_stream_t s);
/* Canonical (b,x,y,z)-lexicographic order of shared/utils.py:116-165 (sort_tensor / sort_points):
 * perm[r] = row holding lexicographic rank r. */
int64_t pcc_lex_order_workspace_bytes(int64_t n);
int pcc_lex_order(const uint64_t* keys, int64_t n, int32_t* perm, void* ws, int64_t ws_bytes,
                  pcc_stream_t s);
/* Same order with fewer radix passes when the caller knows that every coordinate is a multiple of 2^log2_stride (the
 * set's tensor stride) and that frame indices are below 2^frame_bits. */
int pcc_lex_order_bits(const uint64_t* keys, int64_t n, int log2_stride, int frame_bits, int32_t* perm, void* ws,
                       int64_t ws_bytes, pcc_stream_t s);

/* Lossless coordinate coder ("OCT2" streams): fills the slot the reference fills
 * with gpcc_encode / gpcc_decode -- a G-PCC tmc3 subprocess over temporary PLY files
 * (shared/utils.py:169-240, called at codec_pipeline.py:441-462 and codec_parallel.py:266-289) -- for
 * the stride-8 latent coordinates, without leaving the device.  A frame's rows are a sorted run of
 * Morton keys, so the breadth-first octree occupancy bytes follow from prefix comparisons.
 * encode: frame b = rows [frame_offsets_dev[b], frame_offsets_dev[b+1]); its stream is written at
 * out + out_offsets_dev[b] (multiples of 4; reserve pcc_octree_max_bytes(rows of b)) and its size
 * to sizes_dev[b].  Stream: "OCT2" | u32 n | u32 depth | u64 root | u32 nodes[depth] | level bytes.
 * decode: frame b's stream is streams + stream_offsets_dev[b] (multiples of 4), stream_bytes_dev[b]
 * long; its keys (ascending: no sort or de-duplication needed afterwards) go to
 * keys_out[row_offsets_dev[b] .. row_offsets_dev[b+1]), the row counts being the n fields of the
 * stream headers, which the host reads from the bitstream.  status_dev: 0 ok, bit 0 malformed or
 * truncated header, bit 1 inconsistent occupancy bytes (keys of such a frame are set to b << 48). */
int64_t pcc_octree_max_bytes(int64_t n);
int64_t pcc_octree_encode_workspace_bytes(int64_t n_total, int32_t num_frames);
int pcc_octree_encode(const uint64_t* keys, int64_t n_total, const int32_t* frame_offsets_dev,
                      int32_t num_frames, int log2_stride, uint8_t* out,
                      const int64_t* out_offsets_dev, int32_t* sizes_dev, void* ws, int64_t ws_bytes,
                      pcc_stream_t s);
int64_t pcc_octree_decode_workspace_bytes(int64_t n_total);
int pcc_octree_decode(const uint8_t* streams, const int64_t* stream_offsets_dev,
                      const int32_t* stream_bytes_dev, const int32_t* row_offsets_dev,
                      int32_t num_frames, int64_t n_total, int log2_stride, uint64_t* keys_out,
                      int32_t* status_dev, void* ws, int64_t ws_bytes, pcc_stream_t s);

/* ------------------------------------------------------------------------------------------
 * (2) sparse convolution
 * ------------------------------------------------------------------------------------------ */

/* Weight panel packing for pcc_spconv_umma: w [kvol, cin, cout] float32 (ME layout) ->
 * bf16 panels in the UMMA K-major core-matrix layout, cin zero-padded to cin_pad. Host function. */
int64_t pcc_spconv_packed_weight_bytes(int kvol, int cin_pad, int cout);
int pcc_spconv_pack_weights(const float* w, int kvol, int cin, int cout, int cin_pad,
                            void* packed_host);

/* Output-stationary gather-GEMM on tcgen05 (MinkowskiConvolution / ConvolutionTranspose forward
 * inside g_a, h_a, h_s, g_s).  out[o, col0 + n] = act(bias[n] + sum_j in[nbr[j][o]] @ W[j][:, n]
 * (+ residual[o, n])).  nbr == NULL means kvol == 1 and identity rows (1x1 conv and the dense
 * form of the k2/s2 generative transpose).  cin in {16,32,64,128,256} (pad channels with zeros),
 * cout multiple of 16 in [16,256].  ldo / ldr = row strides (elements) of out / residual; either
 * output pointer may be NULL.  cstore: 0 or cout = store all columns; 1..15 = narrow head
 * (occupancy logits C->1, colour head C->3: weights zero-padded to cout = 16, only the first
 * cstore columns are written, ldo >= cstore).
 * nbr is [kvol][nbr_ld] (row pitch nbr_ld >= n_out elements); with the pitch padded to a multiple
 * of 128 (and of 4) each tile's map rows are fetched by one 512-byte cp.async.bulk per offset,
 * entries at columns >= n_out are never used for stores and may hold anything.
 * Kernel: persistent (one CTA per SM), warp specialised (4 epilogue warps, 1 MMA warp, 1 kernel-map
 * loader warp, G groups of 4 gather warps each owning one ring slot); a stage is one kernel offset
 * of a super-tile of T x 128 rows, gathered with cp.async; the MMAs of a stage are interleaved over
 * the T tiles' TMEM accumulators (2 x T accumulators, double buffered against the epilogue); all
 * hand-offs by mbarrier. */
int pcc_spconv_umma(const void* in_bf16, int64_t n_in, int cin, const int32_t* nbr, int64_t nbr_ld,
                    int64_t n_out, int kvol, const void* w_packed, const float* bias, int cout,
                    const void* residual_bf16, int64_t ldr, int relu, void* out_bf16,
                    float* out_f32, int64_t ldo, int cstore, pcc_stream_t s);
/* The same convolution (k = 3, stride 1, C_in = 32) on a coordinate set made of COMPLETE OCTETS -- the un-pruned
 * output of a generative transposed convolution, where the residual blocks and the occupancy head of every g_s level
 * run (receiver/decoder/codec_parallel.py:469): row 8p + o is child o = ox + 2 oy + 4 oz of parent p (pcc_coords_children).
 * Takes the kernel map of the PARENTS (pcc_kmap_k3_blocks on the parent keys at their stride: [27][nbr_ld], nbr_ld a
 * multiple of 4 and >= n_parents rounded up to 128) instead of the 8x larger map of the children; in / out / residual
 * have 8 n_parents rows.  cout = 32, or 16 for a narrow head (cstore < 16 columns stored, like pcc_spconv_umma).
 * Weights are packed by pcc_spconv_octet_pack_weights from the ME layout w [27, cin, cout_src] (columns cout_src..cout
 * zero).  Results equal pcc_spconv_umma's on the children's own kernel map up to fp32 summation order. */
int64_t pcc_spconv_octet_packed_weight_bytes(int cin, int cout);
int pcc_spconv_octet_pack_weights(const float* w, int cin, int cout_src, int cout, void* packed_host);
int pcc_spconv_octet(const void* in_bf16, int64_t n_parents, int cin, const int32_t* nbr_parents, int64_t nbr_ld,
                     const void* w_packed, const float* bias, int cout, const void* residual_bf16, int64_t ldr, int relu,
                     void* out_bf16, float* out_f32, int64_t ldo, int cstore, pcc_stream_t s);

/* Tuning knobs of pcc_spconv_umma (process wide; defaults 0, 4, 0): number of gather groups = ring
 * slots (0 = chosen from the channel count, else 1, 2, 4 or 5) and the maximum number T of 128-row
 * tiles gathered per stage and accumulated concurrently (1, 2 or 4).  use_cg must be 0 (the
 * cp.async.cg gather variant was measured slower and has been removed). */
int pcc_spconv_set_tuning(int groups, int max_tiles, int use_cg);
/* CUDA-core sparse convolution: independent comparator of the tensor-core kernel in the tests
 * (any cin / cout); the product path only uses it for the 1->1 coordinate-only g_s.down_conv.  w is plain bf16 [kvol, cin, cout]. */
int pcc_spconv_simt(const void* in_bf16, int64_t n_in, int cin, const int32_t* nbr, int64_t n_out,
                    int kvol, const void* w_bf16, const float* bias, int cout,
                    const void* residual_bf16, int64_t ldr, int relu, void* out_bf16,
                    float* out_f32, int64_t ldo, pcc_stream_t s);

/* Row utilities. */
int pcc_rows_f32_to_bf16(const float* src, int c_src, const int32_t* src_row, int64_t n_out,
                         void* dst_bf16, int c_dst, pcc_stream_t s);
int pcc_rows_bf16_to_f32(const void* src_bf16, int c_src, int64_t n, float* dst, int c_dst,
                         pcc_stream_t s);
int pcc_gather_rows_bf16(const void* src_bf16, int c, const int32_t* src_row, int64_t n_out,
                         void* dst_bf16, pcc_stream_t s);
int pcc_gather_rows_f32(const float* src, int c, const int32_t* src_row, int64_t n_out, float* dst,
                        pcc_stream_t s);
int pcc_gather_keys(const uint64_t* src, const int32_t* src_row, int64_t n_out, uint64_t* dst,
                    pcc_stream_t s);
/* Prune applied in place for a convolution over the UN-pruned candidates: rows of feats [n, c] (bf16) whose number is not
 * in keep (ascending, m entries -- what pcc_topk_select emits) become zero.  MinkowskiPruning followed by a k3
 * convolution on the kept set (the colour head after the last prune of g_s, receiver/decoder/codec_parallel.py:469)
 * equals, on the kept rows, the same convolution over all candidates with the pruned rows zeroed -- and the candidates
 * are complete octets, so it runs through pcc_spconv_octet on the parents' kernel map and needs no map of its own. */
int pcc_zero_unkept_rows(const int32_t* keep, int64_t m, int64_t n, void* feats_bf16, int c, pcc_stream_t s);

/* Per-frame top-k occupancy selection + pruning (torch.topk + MinkowskiPruning inside
 * g_s(y_hat, k=ks), codec_parallel.py:469).  Keeps, in frame b, the k[b] rows with the largest
 * logits; ties -> lower row (= lower key); NaN ranks last.  keep_row = kept rows ascending,
 * n_keep_dev = count.  Radix select (four 8-bit histogram passes per frame), then one 64-bit scan of
 * (above threshold | tied << 32) marks gives every kept row its output position; num_frames <= 4096.
 * `keys` is only used to document the row order. */
int64_t pcc_topk_workspace_bytes(int64_t n);
int pcc_topk_select(const float* logits, const uint64_t* keys, int64_t n,
                    const int32_t* frame_offsets, const int32_t* k, int32_t num_frames,
                    int32_t* keep_row, int32_t* n_keep_dev, void* ws, int64_t ws_bytes,
                    pcc_stream_t s);

/* ------------------------------------------------------------------------------------------
 * (3) entropy-model kernels.  Latents are [n, C] row-major (rows in key order); symbols and
 *     indexes are channel-major [C, n] with columns in canonical lexicographic order (perm),
 *     the serialisation order of CompressAI (`y.F.t()`, codec_pipeline.py:427).
 * ------------------------------------------------------------------------------------------ */

/* EntropyBottleneck.compress quantisation + the encoder-side decompress round trip
 * (codec_pipeline.py:305-306): sym = round(z - median_c); z_hat = sym + median_c. */
int pcc_factorized_quant(const float* z, const int32_t* perm, const float* medians, int64_t n,
                         int c, int32_t* sym, void* zhat_bf16, float* zhat_f32, pcc_stream_t s);
/* EntropyBottleneck.decompress dequantisation (codec_parallel.py:307). */
int pcc_factorized_dequant(const int32_t* sym, const int32_t* perm, const float* medians,
                           int64_t n, int c, void* zhat_bf16, float* zhat_f32, pcc_stream_t s);
/* Encoder: features_at_coordinates gather + scale gain + build_indexes + quantise for Q qualities
 * at once (codec_pipeline.py:400-430).  params rows are [scales(C) | means(C)]; prow[i] = params
 * row of latent row i (-1 -> zeros); scale is [Q, C]; sym / idx are [Q, C, n]. */
int pcc_gauss_encode(const float* y, const float* params, const int32_t* prow, const int32_t* perm,
                     const float* scale, const float* scale_table, int table_len, float scale_bound,
                     int64_t n, int c, int q, int32_t* sym, uint16_t* idx, pcc_stream_t s);
/* Decoder: indexes only (codec_parallel.py:386-398). */
int pcc_gauss_indexes(const float* params, const int32_t* prow, const int32_t* perm,
                      const float* scale, const float* scale_table, int table_len,
                      float scale_bound, int64_t n, int c, uint16_t* idx, pcc_stream_t s);
/* Decoder: dead-zone offset dequantisation (codec_parallel.py:401-409):
 * y_hat = sign(q) * (|q| - off(max(sigma*s, bound), s)) / s + mu, off = 0 where |q| < 1e-4. */
int pcc_gauss_dequant(const int32_t* sym, const float* params, const int32_t* prow,
                      const int32_t* perm, const float* scale, const float* offset_gain,
                      float scale_bound, int64_t n, int c, void* yhat_bf16, float* yhat_f32,
                      pcc_stream_t s);

/* CompressAI _CXX.pmf_to_quantized_cdf (host; used by update(), codec_pipeline.py:69).
 * cdf_out has pmf_len + 1 entries. Returns PCC_E_INVALID if the pmf cannot be quantised. */
int pcc_pmf_to_quantized_cdf(const float* pmf, int pmf_len, int precision, int32_t* cdf_out);

/* ------------------------------------------------------------------------------------------
 * (4) interleaved multi-lane rANS ("PCL1" container)
 *     Symbol semantics follow CompressAI's encode_with_indexes / decode_with_indexes
 *     (escape to 4-bit bypass nibbles for values outside the table support), so any int32
 *     symbol round-trips.  One warp = one segment of seg_size symbols = 32 interleaved lanes.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  const uint32_t* cdf;      /* flat quantised CDFs, 16-bit precision (last entry of a row = 65536) */
  const int32_t* cdf_off;   /* [T] start of table t in cdf */
  const int32_t* cdf_size;  /* [T] entries in table t (= pmf_length + 2) */
  const int32_t* offset;    /* [T] symbol offset of table t */
  const uint16_t* lut;      /* [T][257] decode search windows from pcc_rans_build_lut, or NULL */
  int32_t num_tables;
  int32_t cdf_total;        /* entries in cdf over all tables; 0 = unknown (decoder then reads the
                               tables from global memory instead of staging them in shared memory) */
} pcc_cdf_tables_t;

/* Decoder acceleration table: lut[t][b] = largest s with cdf_t[s] <= b << 8 (b = 0..255),
 * lut[t][256] = last symbol; narrows the CDF search to a few entries.  pcc_rans_lut_entries()
 * uint16 elements; build once per table set (model.update()). */
int64_t pcc_rans_lut_entries(int32_t num_tables);
int pcc_rans_build_lut(const pcc_cdf_tables_t* tables, uint16_t* lut, pcc_stream_t s);

/* Upper bound on container bytes for n symbols. */
int64_t pcc_rans_max_bytes(int64_t n, int32_t seg_size);
int64_t pcc_rans_workspace_bytes(int64_t n, int32_t seg_size);
/* idx == NULL: table index = i / chan_len (EntropyBottleneck: index = channel).
 * out: device buffer of pcc_rans_max_bytes; out_bytes_dev: total container size. */
int pcc_rans_encode(const int32_t* sym, const uint16_t* idx, int64_t chan_len, int64_t n,
                    int32_t seg_size, const pcc_cdf_tables_t* tables, uint8_t* out,
                    int32_t* out_bytes_dev, void* ws, int64_t ws_bytes, pcc_stream_t s);
/* All quality levels of a GOP in one call (the reference codes them in ONE batched
 * gaussian_conditional.compress call: sender/encoder/codec_pipeline.py:414-432).  sym / idx: [num_streams][n];
 * stream q becomes an independent PCL1 container at out + q * out_stride (out_stride: multiple of 4,
 * >= pcc_rans_max_bytes(n, seg_size)), byte-identical to pcc_rans_encode of that stream alone; its size lands in
 * out_bytes_dev[q].  Three launches whatever num_streams is (grid.y = stream): 25 levels x 58 segments fill the
 * machine that one stream leaves idle.  Workspace: num_streams * pcc_rans_workspace_bytes(n, seg_size). */
int pcc_rans_encode_batch(const int32_t* sym, const uint16_t* idx, int64_t chan_len, int64_t n, int32_t num_streams,
                          int32_t seg_size, const pcc_cdf_tables_t* tables, uint8_t* out, int64_t out_stride,
                          int32_t* out_bytes_dev, void* ws, int64_t ws_bytes, pcc_stream_t s);
/* seg_size = the container header's value (the caller holds the host copy of the bytes).
 * status_dev: 0 ok, 1 truncated container, 2 corrupt stream, 3 header mismatch. */
int pcc_rans_decode(const uint8_t* container, int64_t container_bytes, int32_t seg_size,
                    const uint16_t* idx, int64_t chan_len, int64_t n,
                    const pcc_cdf_tables_t* tables, int32_t* sym, int32_t* status_dev,
                    pcc_stream_t s);

/* ------------------------------------------------------------------------------------------
 * (5) launch sequencer: a whole transform as one call (replaces the per-layer
 *     Python dispatch of codec_pipeline.py:273-354 / codec_parallel.py:302-469 and the thread-per-
 *     stage queues of codec_pipeline.py:31-54).  Entry i runs the function named by `op` with the
 *     arguments a[0..] in that function's own order (pointers as integers, stream omitted), all on
 *     stream s.  Stops at the first failing entry (its index in *failed_op, else -1).
 * ------------------------------------------------------------------------------------------ */
#define PCC_OP_SPCONV_UMMA 1
#define PCC_OP_KMAP_DOWN2 2
#define PCC_OP_BLOCKMAP_BUILD 3
#define PCC_OP_KMAP_K3_BLOCKS 4
#define PCC_OP_ROWS_F32_TO_BF16 5
#define PCC_OP_COORDS_CHILDREN 6
#define PCC_OP_GATHER_ROWS_BF16 7
#define PCC_OP_GATHER_KEYS 8
#define PCC_OP_FRAME_OFFSETS 9
#define PCC_OP_TOPK_SELECT 10
#define PCC_OP_SPCONV_OCTET 11
#define PCC_OP_ZERO_UNKEPT_ROWS 12
#define PCC_OP_GATHER_ROWS_F32 13
typedef struct {
  int64_t op;
  int64_t a[19];
} pcc_op_t;
int pcc_run_program(const pcc_op_t* ops, int32_t n_ops, pcc_stream_t s, int32_t* failed_op);

/* ------------------------------------------------------------------------------------------
 * (6) bitstream containers of the reference, host side only: no GPU, no
 *     allocation.  Field order, widths and endianness are the reference's (int32 / float64, big endian).
 *     Batched layout  = make_bitstream_batched / read_bitstream_batched
 *                       (sender/encoder/codec_pipeline.py:464-517, receiver/decoder/codec_parallel.py:173-216);
 *     per-frame layout = make_bitstream / read_bitstream
 *                       (sender/encoder/codec_single.py:269-337, receiver/decoder/codec_single.py:90-136).
 *     The y/z strings are PCL1 lane containers (4), points_i is the OCT2 stream of (1b).
 *     ks holds three rows of per-frame point counts: ks[level * ks_ld + frame], level 0..2.
 *     The parse functions return spans that point INTO `data`; PCC_E_INVALID for a truncated or
 *     inconsistent blob (negative length, more than max_frames frames).
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  const uint8_t* data;
  int64_t len;
} pcc_span_t;
int64_t pcc_container_batched_bytes(int32_t num_frames, int64_t y_len, int64_t z_len, const pcc_span_t* points);
int pcc_container_write_batched(int32_t num_frames, double q_geo, double q_att, int32_t y_shape, int32_t z_shape,
                                pcc_span_t y, pcc_span_t z, const pcc_span_t* points, const int32_t* ks,
                                int32_t ks_ld, uint8_t* out, int64_t out_cap, int64_t* written);
int pcc_container_parse_batched(const uint8_t* data, int64_t nbytes, int32_t max_frames, int32_t* num_frames,
                                double* q_geo, double* q_att, int32_t* y_shape, int32_t* z_shape, pcc_span_t* y,
                                pcc_span_t* z, pcc_span_t* points, int32_t* ks, int32_t ks_ld);
int64_t pcc_container_frames_bytes(int32_t num_frames, const pcc_span_t* y, const pcc_span_t* z,
                                   const pcc_span_t* points);
int pcc_container_write_frames(int32_t num_frames, double q_geo, double q_att, const int32_t* y_shapes,
                               const int32_t* z_shapes, const pcc_span_t* y, const pcc_span_t* z,
                               const pcc_span_t* points, const int32_t* ks, int32_t ks_ld, uint8_t* out,
                               int64_t out_cap, int64_t* written);
int pcc_container_parse_frames(const uint8_t* data, int64_t nbytes, int32_t max_frames, int32_t* num_frames,
                               double* q_geo, double* q_att, int32_t* y_shapes, int32_t* z_shapes, pcc_span_t* y,
                               pcc_span_t* z, pcc_span_t* points, int32_t* ks, int32_t ks_ld);

/* ------------------------------------------------------------------------------------------
 * (7) host -> device staging of a raw frame (replaces torch.tensor(numpy, device=) of
 *     sender/encoder/codec_pipeline.py:252-266, a pageable synchronous copy).  The n_src host arrays are
 *     laid back to back; a pool of `threads` native threads (the caller is one of them) copies them into
 *     `pinned` (page-locked, same layout) in pieces of piece_bytes and enqueues each piece's DMA to `dev`
 *     on stream s as soon as that piece is copied.  Returns when every DMA is enqueued; the data is on
 *     the device when s reaches that point.  Safe to call from several host threads.
 *     convert = PCC_STAGE_F64_TO_BF16: the sources are float64 arrays and every value is converted on the way
 *     (double -> float -> bf16, both round to nearest even: the roundings pcc_ingest_features applies), so
 *     a colour travels as 2 bytes instead of 8; pinned / dev then hold nbytes / 4 per source, back to back.
 * ------------------------------------------------------------------------------------------ */
#define PCC_STAGE_COPY 0
#define PCC_STAGE_F64_TO_BF16 1
int pcc_stage_h2d(const void* const* srcs, const int64_t* nbytes, int32_t n_src, void* pinned, void* dev,
                  int64_t piece_bytes, int32_t threads, int32_t convert, pcc_stream_t s);
/* Host helpers behind pcc_stage_h2d (exported for tests and C hosts that stage by themselves): the
 * conversion of PCC_STAGE_F64_TO_BF16, and a memcpy with cache-bypassing stores (the DMA engine reads the
 * pinned buffer next; lines left dirty in the CPU caches slow that read down). */
void pcc_host_f64_to_bf16(const double* src, int64_t n, uint16_t* dst);
void pcc_host_copy_stream(const void* src, int64_t nbytes, void* dst);

/* ------------------------------------------------------------------------------------------
 * (8) all byte strings of a GOP in one device buffer, so the host fetches them with a single copy
 *     (the reference brings every string to the host on its own: codec_pipeline.py:305-306, 426-432,
 *     441-462).  bufs[i] / lens[i] are device pointers to string i and to its int32 length counter (host
 *     arrays of n pointers).  out = u32 len[n] | u32 end[ceil(n / 64)] | pad to 16 | payloads, each
 *     starting on a 16-byte boundary, string 0 at pcc_pack_strings_header_bytes(n); end[last] is the
 *     number of bytes used.  A string that does not fit in out_cap is left out (its length is still
 *     written and end[] still advances: the caller compares end with what it copied).
 * ------------------------------------------------------------------------------------------ */
int64_t pcc_pack_strings_header_bytes(int32_t n);
int pcc_pack_strings(const void* const* bufs, const int32_t* const* lens, int32_t n, uint8_t* out,
                     int64_t out_cap, pcc_stream_t s);

/* ------------------------------------------------------------------------------------------
 * (9) capturer frame preparation on the device (capture half; replaces the numpy +
 *     Open3D body of Capturer.get_zed_frames, sender/capturer/capturer.py:84-120).  xyzrgba = the camera's
 *     XYZRGBA measure, float32 [n_px, 4] on the device (16-byte aligned; the colour float carries r, g, b in
 *     its three low bytes, :87-90).  Pixels that are not finite or farther than depth_clip (float32 norm,
 *     :93-94) are dropped; the rest is voxel-down-sampled as Open3D does it in float64 (:99-106: bound =
 *     min - voxel_size / 2, index = floor((p - bound) / voxel_size), running sums in input order, mean),
 *     rounded half-to-even to the voxel grid as int16 (:107), de-duplicated keeping the first row in
 *     (x,y,z) order (:110-112) and, when max_points > 0 and more rows are left, cut to the max_points rows of
 *     largest z (:117-120).  Where the reference leaves an order to its libraries (unordered_map iteration,
 *     argpartition ties) this follows oracle/capture_ref.py: voxels in ascending index order, z ties to the
 *     earlier row, output in (x,y,z) order.
 *     Output: points int16 [m,3], colors float64 [m,3] (capacity rows each), *n_out_dev = m, or -1 when a
 *     voxel index needs more than 21 bits, -2 when a rounded coordinate leaves int16; rows beyond capacity
 *     are not written (m is still reported).  No host synchronisation.
 * ------------------------------------------------------------------------------------------ */
int64_t pcc_capture_voxelize_workspace_bytes(int64_t n_px);
int pcc_capture_voxelize(const float* xyzrgba, int64_t n_px, double voxel_size, float depth_clip,
                         int32_t max_points, int16_t* points, double* colors, int64_t capacity,
                         int32_t* n_out_dev, void* ws, int64_t ws_bytes, pcc_stream_t s);

#ifdef __cplusplus
}
#endif
#endif /* PCC_B200_H */
