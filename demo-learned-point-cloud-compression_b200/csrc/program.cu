// Launch sequencer: one C call runs a list of operations of this library on one stream.
//
// The reference drives every sparse convolution from Python (one MinkowskiEngine call per layer,
// sender/encoder/codec_pipeline.py:273-354, receiver/decoder/codec_parallel.py:302-469), one thread per
// stage (codec_pipeline.py:31-54).  Here a whole transform is a short program of ~40 entries whose
// buffers are raw device addresses inside a caller-owned arena: the host cost per launch drops from a
// Python call (tensor allocation + argument marshalling, ~12 us) to a table walk.
// Sizes differ from frame to frame, so the program is rebuilt per call; nothing is captured or cached.
#include "common.cuh"

using namespace pcc;

extern "C" {

int pcc_run_program(const pcc_op_t* ops, int32_t n_ops, pcc_stream_t s, int32_t* failed_op) {
  PCC_CHECK_ARG(n_ops >= 0 && (ops || n_ops == 0));
  for (int32_t i = 0; i < n_ops; ++i) {
    const int64_t* a = ops[i].a;
    int rc = PCC_OK;
    switch (ops[i].op) {
      case PCC_OP_SPCONV_UMMA:
        rc = pcc_spconv_umma((const void*)a[0], a[1], (int)a[2], (const int32_t*)a[3], a[4], a[5], (int)a[6],
                             (const void*)a[7], (const float*)a[8], (int)a[9], (const void*)a[10], a[11], (int)a[12],
                             (void*)a[13], (float*)a[14], a[15], (int)a[16], s);
        break;
      case PCC_OP_KMAP_DOWN2:
        rc = pcc_kmap_down2((const uint64_t*)a[0], (const int32_t*)a[1], a[2], (int)a[3], (int32_t*)a[4], a[5], a[6], s);
        break;
      case PCC_OP_BLOCKMAP_BUILD:
        rc = pcc_blockmap_build((const uint64_t*)a[0], a[1], (int)a[2], (uint64_t*)a[3], (void*)a[4], a[5], s);
        break;
      case PCC_OP_KMAP_K3_BLOCKS:
        rc = pcc_kmap_k3_blocks((const uint64_t*)a[0], a[1], (int)a[2], (const uint64_t*)a[3], (const void*)a[4], a[5],
                                (int32_t*)a[6], a[7], s);
        break;
      case PCC_OP_ROWS_F32_TO_BF16:
        rc = pcc_rows_f32_to_bf16((const float*)a[0], (int)a[1], (const int32_t*)a[2], a[3], (void*)a[4], (int)a[5], s);
        break;
      case PCC_OP_COORDS_CHILDREN:
        rc = pcc_coords_children((const uint64_t*)a[0], a[1], (int)a[2], (uint64_t*)a[3], s);
        break;
      case PCC_OP_GATHER_ROWS_BF16:
        rc = pcc_gather_rows_bf16((const void*)a[0], (int)a[1], (const int32_t*)a[2], a[3], (void*)a[4], s);
        break;
      case PCC_OP_GATHER_KEYS:
        rc = pcc_gather_keys((const uint64_t*)a[0], (const int32_t*)a[1], a[2], (uint64_t*)a[3], s);
        break;
      case PCC_OP_FRAME_OFFSETS:
        rc = pcc_frame_offsets((const uint64_t*)a[0], a[1], (int32_t)a[2], (int32_t*)a[3], s);
        break;
      case PCC_OP_TOPK_SELECT:
        rc = pcc_topk_select((const float*)a[0], (const uint64_t*)a[1], a[2], (const int32_t*)a[3], (const int32_t*)a[4],
                             (int32_t)a[5], (int32_t*)a[6], (int32_t*)a[7], (void*)a[8], a[9], s);
        break;
      case PCC_OP_SPCONV_OCTET:
        rc = pcc_spconv_octet((const void*)a[0], a[1], (int)a[2], (const int32_t*)a[3], a[4], (const void*)a[5],
                              (const float*)a[6], (int)a[7], (const void*)a[8], a[9], (int)a[10], (void*)a[11],
                              (float*)a[12], a[13], (int)a[14], s);
        break;
      case PCC_OP_ZERO_UNKEPT_ROWS:
        rc = pcc_zero_unkept_rows((const int32_t*)a[0], a[1], a[2], (void*)a[3], (int)a[4], s);
        break;
      case PCC_OP_GATHER_ROWS_F32:
        rc = pcc_gather_rows_f32((const float*)a[0], (int)a[1], (const int32_t*)a[2], a[3], (float*)a[4], s);
        break;
      default:
        set_error("pcc_run_program: unknown op %lld at entry %d", (long long)ops[i].op, (int)i);
        rc = PCC_E_INVALID;
    }
    if (rc != PCC_OK) {
      if (failed_op) *failed_op = i;
      return rc;
    }
  }
  if (failed_op) *failed_op = -1;
  return PCC_OK;
}

}  // extern "C"
