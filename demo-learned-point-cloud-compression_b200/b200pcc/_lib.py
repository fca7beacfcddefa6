"""ctypes binding of libpcc_b200.so (C ABI declared in include/pcc_b200.h).

The library is the product: importing this module fails loudly when the shared object has not been
built (``python -c "import __graft_entry__ as g; g.build()"`` or ``make -C csrc``).  There is no CPU
fallback anywhere in the package.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpcc_b200.so")


class PccError(RuntimeError):
    pass


if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build the CUDA library first (make -C "
        f"{os.path.join(os.path.dirname(_HERE), 'csrc')}). b200pcc has no CPU fallback."
    )

lib = C.CDLL(LIB_PATH)

P = C.c_void_p
I32 = C.c_int32
I64 = C.c_int64
INT = C.c_int
F32 = C.c_float


class Span(C.Structure):
    _fields_ = [("data", P), ("len", I64)]


class CdfTables(C.Structure):
    _fields_ = [("cdf", P), ("cdf_off", P), ("cdf_size", P), ("offset", P), ("lut", P), ("num_tables", I32), ("cdf_total", I32)]


# name -> (restype, argtypes); must list every symbol include/pcc_b200.h declares
SIGNATURES = {
    "pcc_version": (INT, []),
    "pcc_last_error": (C.c_char_p, []),
    "pcc_launch_count": (I64, []),
    "pcc_kmap_count_pairs": (INT, [P, I64, I64, INT, P, P]),
    "pcc_fault_fetch": (INT, [P, P]),
    "pcc_fetch_words": (INT, [P, I32, I32, P, P]),
    "pcc_fault_string": (C.c_char_p, [I32]),
    "pcc_debug_inject": (INT, [I32]),
    "pcc_debug_trace_words": (I64, []),
    "pcc_debug_trace": (INT, [P]),
    "pcc_keys_from_coords_i32": (INT, [P, I64, P, P]),
    "pcc_keys_from_coords_f32": (INT, [P, I64, P, P]),
    "pcc_ingest_frame": (INT, [P, P, I64, I32, P, P, P]),
    "pcc_ingest_points": (INT, [P, I64, I32, P, P]),
    "pcc_ingest_features": (INT, [P, P, I64, P, INT, P]),
    "pcc_ingest_features_bf16": (INT, [P, P, I64, P, INT, P]),
    "pcc_unpack_output": (INT, [P, P, I64, P, P, P]),
    "pcc_pack_visualizer": (INT, [P, P, I64, F32, P, P, P]),
    "pcc_coords_from_keys": (INT, [P, I64, P, P]),
    "pcc_sort_unique_workspace_bytes": (I64, [I64]),
    "pcc_sort_unique": (INT, [P, I64, P, P, P, P, I64, P]),
    "pcc_sort_unique_bits": (INT, [P, I64, INT, P, P, P, P, I64, P]),
    "pcc_coords_downsample_workspace_bytes": (I64, [I64]),
    "pcc_coords_downsample": (INT, [P, I64, INT, P, P, P, P, I64, P]),
    "pcc_coords_downsample_static_max_rows": (I64, []),
    "pcc_coords_downsample_static": (INT, [P, I64, INT, P, P, P, P]),
    "pcc_coords_pyramid_workspace_bytes": (I64, [I64]),
    "pcc_coords_pyramid": (INT, [P, I64, INT, INT, P, P, P, I32, P, P, I64, P]),
    "pcc_coords_children": (INT, [P, I64, INT, P, P]),
    "pcc_hash_build": (INT, [P, I64, P, P, I64, P]),
    "pcc_hash_query": (INT, [P, I64, P, P, I64, P, P]),
    "pcc_kmap_k3": (INT, [P, I64, INT, P, P, I64, P, P]),
    "pcc_blockmap_build": (INT, [P, I64, INT, P, P, I64, P]),
    "pcc_kmap_k3_blocks": (INT, [P, I64, INT, P, P, I64, P, I64, P]),
    "pcc_kmap_down2": (INT, [P, P, I64, INT, P, I64, I64, P]),
    "pcc_frame_offsets": (INT, [P, I64, I32, P, P]),
    "pcc_lex_order_workspace_bytes": (I64, [I64]),
    "pcc_lex_order": (INT, [P, I64, P, P, I64, P]),
    "pcc_lex_order_bits": (INT, [P, I64, INT, INT, P, P, I64, P]),
    "pcc_octree_max_bytes": (I64, [I64]),
    "pcc_octree_encode_workspace_bytes": (I64, [I64, I32]),
    "pcc_octree_encode": (INT, [P, I64, P, I32, INT, P, P, P, P, I64, P]),
    "pcc_octree_decode_workspace_bytes": (I64, [I64]),
    "pcc_octree_decode": (INT, [P, P, P, P, I32, I64, INT, P, P, P, I64, P]),
    "pcc_spconv_packed_weight_bytes": (I64, [INT, INT, INT]),
    "pcc_spconv_pack_weights": (INT, [P, INT, INT, INT, INT, P]),
    "pcc_spconv_umma": (INT, [P, I64, INT, P, I64, I64, INT, P, P, INT, P, I64, INT, P, P, I64, INT, P]),
    "pcc_spconv_octet_packed_weight_bytes": (I64, [INT, INT]),
    "pcc_spconv_octet_pack_weights": (INT, [P, INT, INT, INT, P]),
    "pcc_spconv_octet": (INT, [P, I64, INT, P, I64, P, P, INT, P, I64, INT, P, P, I64, INT, P]),
    "pcc_spconv_set_tuning": (INT, [INT, INT, INT]),
    "pcc_spconv_simt": (INT, [P, I64, INT, P, I64, INT, P, P, INT, P, I64, INT, P, P, I64, P]),
    "pcc_rows_f32_to_bf16": (INT, [P, INT, P, I64, P, INT, P]),
    "pcc_rows_bf16_to_f32": (INT, [P, INT, I64, P, INT, P]),
    "pcc_gather_rows_bf16": (INT, [P, INT, P, I64, P, P]),
    "pcc_gather_rows_f32": (INT, [P, INT, P, I64, P, P]),
    "pcc_zero_unkept_rows": (INT, [P, I64, I64, P, INT, P]),
    "pcc_gather_keys": (INT, [P, P, I64, P, P]),
    "pcc_topk_workspace_bytes": (I64, [I64]),
    "pcc_topk_select": (INT, [P, P, I64, P, P, I32, P, P, P, I64, P]),
    "pcc_factorized_quant": (INT, [P, P, P, I64, INT, P, P, P, P]),
    "pcc_factorized_dequant": (INT, [P, P, P, I64, INT, P, P, P]),
    "pcc_gauss_encode": (INT, [P, P, P, P, P, P, INT, F32, I64, INT, INT, P, P, P]),
    "pcc_gauss_indexes": (INT, [P, P, P, P, P, INT, F32, I64, INT, P, P]),
    "pcc_gauss_dequant": (INT, [P, P, P, P, P, P, F32, I64, INT, P, P, P]),
    "pcc_pmf_to_quantized_cdf": (INT, [P, INT, INT, P]),
    "pcc_rans_max_bytes": (I64, [I64, I32]),
    "pcc_rans_workspace_bytes": (I64, [I64, I32]),
    "pcc_rans_lut_entries": (I64, [I32]),
    "pcc_rans_build_lut": (INT, [C.POINTER(CdfTables), P, P]),
    "pcc_rans_encode": (INT, [P, P, I64, I64, I32, C.POINTER(CdfTables), P, P, P, I64, P]),
    "pcc_rans_encode_batch": (INT, [P, P, I64, I64, I32, I32, C.POINTER(CdfTables), P, I64, P, P, I64, P]),
    "pcc_rans_decode": (INT, [P, I64, I32, P, I64, I64, C.POINTER(CdfTables), P, P, P]),
    "pcc_run_program": (INT, [P, I32, P, P]),
    "pcc_container_batched_bytes": (I64, [I32, I64, I64, P]),
    "pcc_container_write_batched": (INT, [I32, C.c_double, C.c_double, I32, I32, Span, Span, P, P, I32, P, I64, P]),
    "pcc_container_parse_batched": (INT, [P, I64, I32, P, P, P, P, P, P, P, P, P, I32]),
    "pcc_container_frames_bytes": (I64, [I32, P, P, P]),
    "pcc_container_write_frames": (INT, [I32, C.c_double, C.c_double, P, P, P, P, P, P, I32, P, I64, P]),
    "pcc_container_parse_frames": (INT, [P, I64, I32, P, P, P, P, P, P, P, P, P, I32]),
    "pcc_stage_h2d": (INT, [P, P, I32, P, P, I64, I32, I32, P]),
    "pcc_host_f64_to_bf16": (None, [P, I64, P]),
    "pcc_host_copy_stream": (None, [P, I64, P]),
    "pcc_pack_strings_header_bytes": (I64, [I32]),
    "pcc_pack_strings": (INT, [P, P, I32, P, I64, P]),
    "pcc_capture_voxelize_workspace_bytes": (I64, [I64]),
    "pcc_capture_voxelize": (INT, [P, I64, C.c_double, F32, I32, P, P, I64, P, P, I64, P]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)  # AttributeError here = header/library mismatch: fail loudly
    _fn.restype = _res
    _fn.argtypes = _args


def last_error() -> str:
    return lib.pcc_last_error().decode("utf-8", "replace")


def check(rc: int) -> None:
    if rc != 0:
        raise PccError(f"pcc_b200 error {rc}: {last_error()}")


def launch_count() -> int:
    return int(lib.pcc_launch_count())
