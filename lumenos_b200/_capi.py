"""ctypes declarations of include/lumenos_b200.h.  Loading fails loudly when the CUDA library is not built:
there is no CPU fallback in this package."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("LUMENOS_B200_SO") or os.path.join(_HERE, "liblumenos_b200.so")   # override: A/B builds of tuning variants

u64p = C.POINTER(C.c_uint64)
u32p = C.POINTER(C.c_uint32)
vp = C.c_void_p


class LmnLeafFormat(C.Structure):
    _fields_ = [("header", C.c_char_p), ("header_len", C.c_uint32), ("poly_header", C.c_char_p), ("poly_header_len", C.c_uint32),
                ("limb_header", C.c_char_p), ("limb_header_len", C.c_uint32)]


class LmnLigeroMetadata(C.Structure):
    _fields_ = [("rows", C.c_uint32), ("cols", C.c_uint32), ("rho_inv", C.c_uint32), ("queries", C.c_uint32)]


class LmnParams(C.Structure):
    _fields_ = [("log_n", C.c_uint32), ("num_q", C.c_uint32), ("num_p", C.c_uint32),
                ("q", u64p), ("p", u64p), ("psi_q", u64p), ("psi_p", u64p), ("t", C.c_uint64),
                ("scalar_centered", C.c_uint32), ("moddown_round_nearest", C.c_uint32),
                ("innersum_rowswap_last", C.c_uint32)]


# name -> (restype, argtypes); exactly the symbols include/lumenos_b200.h declares
SIGNATURES = {
    "lmn_ctx_create": (C.c_int, [C.POINTER(LmnParams), C.c_int, C.POINTER(vp)]),
    "lmn_ctx_destroy": (None, [vp]),
    "lmn_last_error": (C.c_char_p, []),
    "lmn_ctx_set_stream": (C.c_int, [vp, vp]),
    "lmn_ctx_synchronize": (C.c_int, [vp]),
    "lmn_ctx_launch_count": (C.c_uint64, [vp]),
    "lmn_ctx_trim": (C.c_int, [vp, C.POINTER(C.c_size_t)]),
    "lmn_ctx_set_batch": (C.c_int, [vp, C.c_uint32]),
    "lmn_store_create": (C.c_int, [vp, C.c_uint32, C.c_uint32, C.POINTER(vp)]),
    "lmn_store_wrap": (C.c_int, [vp, vp, C.c_uint32, C.c_uint32, C.POINTER(vp)]),
    "lmn_store_create_small": (C.c_int, [vp, C.c_uint32, C.POINTER(vp)]),
    "lmn_store_destroy": (None, [vp]),
    "lmn_store_count": (C.c_uint32, [vp]),
    "lmn_store_level": (C.c_uint32, [vp]),
    "lmn_store_device_ptr": (vp, [vp]),
    "lmn_store_bytes": (C.c_size_t, [vp]),
    "lmn_store_upload": (C.c_int, [vp, C.c_uint32, C.c_uint32, vp]),
    "lmn_store_download": (C.c_int, [vp, C.c_uint32, C.c_uint32, vp]),
    "lmn_store_upload_limbs": (C.c_int, [vp, C.c_uint32, C.c_uint32, vp]),
    "lmn_store_download_limbs": (C.c_int, [vp, C.c_uint32, C.c_uint32, vp]),
    "lmn_store_upload_limb": (C.c_int, [vp, C.c_uint32, C.c_uint32, C.c_uint32, vp]),
    "lmn_store_download_limb": (C.c_int, [vp, C.c_uint32, C.c_uint32, C.c_uint32, vp]),
    "lmn_host_alloc": (C.c_int, [C.c_size_t, C.POINTER(vp)]),
    "lmn_host_free": (None, [vp]),
    "lmn_plan_create": (C.c_int, [vp, C.c_uint32, C.c_uint32, vp, C.c_uint32, C.POINTER(vp)]),
    "lmn_plan_destroy": (None, [vp]),
    "lmn_plan_counts": (C.c_int, [vp, u64p, u64p, u64p]),
    "lmn_plan_size": (C.c_uint32, [vp]),
    "lmn_plan_host_check": (C.c_int, [C.c_uint32, vp, C.c_uint64, vp, vp]),
    "lmn_encode": (C.c_int, [vp, vp, vp, vp, C.c_uint32, vp]),
    "lmn_encode_host": (C.c_int, [vp, vp, vp, vp, vp, C.c_uint32, vp]),
    "lmn_inner_product2_host": (C.c_int, [vp, vp, vp, vp, C.c_uint32, vp, vp, vp, vp]),
    "lmn_ntt": (C.c_int, [vp, vp, vp, vp]),
    "lmn_encode_lanes": (C.c_int, [vp, vp, vp, vp, vp, C.c_uint32]),
    "lmn_profile_enable": (C.c_int, [vp, C.c_uint32]),
    "lmn_profile_read": (C.c_int, [vp, C.c_int, C.POINTER(C.c_double), u64p, u64p, C.c_int]),
    "lmn_kernel_count": (C.c_int, []),
    "lmn_kernel_name": (C.c_char_p, [C.c_int]),
    "lmn_bench_int_peak": (C.c_int, [vp, C.POINTER(C.c_double)]),
    "lmn_bench_product_butterfly": (C.c_int, [vp, C.POINTER(C.c_double)]),
    "lmn_comm_unique_id": (C.c_int, [C.c_char_p]),
    "lmn_comm_init": (C.c_int, [vp, C.c_char_p, C.c_int, C.c_int]),
    "lmn_comm_destroy": (C.c_int, [vp]),
    "lmn_comm_rank": (C.c_int, [vp]),
    "lmn_comm_world": (C.c_int, [vp]),
    "lmn_exchange_queried": (C.c_int, [vp, vp, C.c_uint32, vp, C.c_uint32, vp]),
    "lmn_exchange_encoded": (C.c_int, [vp, vp, C.c_uint32, vp]),
    "lmn_commit_leaves_from_slices": (C.c_int, [vp, vp, C.c_uint32, vp]),
    "lmn_exchange_to_slices": (C.c_int, [vp, vp, C.c_uint32, vp]),
    "lmn_allgather_leaf_hashes": (C.c_int, [vp, C.c_char_p, C.c_uint32, C.c_char_p]),
    "lmn_load_plaintext_ring": (C.c_int, [vp, C.c_uint64]),
    "lmn_load_public_key": (C.c_int, [vp, vp]),
    "lmn_load_secret_key": (C.c_int, [vp, vp]),
    "lmn_witness_columns": (C.c_int, [vp, C.c_char_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, vp]),
    "lmn_encrypt_columns": (C.c_int, [vp, vp, C.c_char_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_char_p, vp]),
    "lmn_encrypt_zero": (C.c_int, [vp, C.c_char_p, C.c_uint32, vp, C.c_uint32]),
    "lmn_test_encryption_samples": (C.c_int, [vp, vp, C.c_uint32, C.c_int, vp]),
    "lmn_encode_plaintext": (C.c_int, [vp, vp, C.c_uint32, vp]),
    "lmn_decrypt_decode": (C.c_int, [vp, vp, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint32, vp]),
    "lmn_load_galois_key": (C.c_int, [vp, C.c_uint64, vp]),
    "lmn_load_ring_switch_key": (C.c_int, [vp, C.c_uint32, vp, C.c_uint32, C.c_uint32, C.c_uint64]),
    "lmn_inner_product": (C.c_int, [vp, vp, vp, C.c_uint32, vp]),
    "lmn_inner_product2": (C.c_int, [vp, vp, vp, vp, C.c_uint32, vp, vp]),
    "lmn_ring_switch": (C.c_int, [vp, vp, vp]),
    "lmn_query_columns": (C.c_int, [vp, vp, vp, C.c_uint32, vp]),
    "lmn_commit_leaves": (C.c_int, [vp, vp, vp]),
    "lmn_store_gather": (C.c_int, [vp, vp, vp, C.c_uint32, vp]),
    "lmn_merkle_build": (C.c_int, [vp, vp, C.POINTER(LmnLeafFormat), C.POINTER(vp)]),
    "lmn_leaf_hashes": (C.c_int, [vp, vp, C.POINTER(LmnLeafFormat), vp]),
    "lmn_merkle_from_leaf_hashes": (C.c_int, [vp, vp, C.c_uint32, C.POINTER(vp)]),
    "lmn_merkle_from_leaf_hashes_noalloc": (C.c_int, [vp, vp, C.c_uint32, vp, vp, C.c_size_t]),
    "lmn_merkle_destroy": (None, [vp]),
    "lmn_merkle_leaf_count": (C.c_uint32, [vp]),
    "lmn_merkle_depth": (C.c_uint32, [vp]),
    "lmn_merkle_root": (C.c_int, [vp, vp]),
    "lmn_merkle_leaf_hash": (C.c_int, [vp, C.c_uint32, vp]),
    "lmn_merkle_path": (C.c_int, [vp, C.c_uint32, vp]),
    "lmn_leaf_bytes": (C.c_int, [vp, vp, C.c_uint32, C.POINTER(LmnLeafFormat), vp, C.c_size_t, C.POINTER(C.c_size_t)]),
    "lmn_store_marshal": (C.c_int, [vp, vp, C.c_uint32, C.c_uint32, C.POINTER(LmnLeafFormat), vp, C.c_size_t, C.POINTER(C.c_size_t)]),
    "lmn_proof_marshal": (C.c_int, [vp, C.POINTER(LmnLigeroMetadata), vp, vp, vp, C.POINTER(LmnLeafFormat), C.POINTER(LmnLeafFormat),
                                    C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t, vp, C.c_size_t, C.POINTER(C.c_size_t)]),
    "lmn_rescale_scale": (C.c_uint64, [vp, C.c_uint32, C.c_uint32]),
    "lmn_ct_add": (C.c_int, [vp, vp, C.c_uint32, vp, C.c_uint32, vp, C.c_uint32]),
    "lmn_ct_sub": (C.c_int, [vp, vp, C.c_uint32, vp, C.c_uint32, vp, C.c_uint32]),
    "lmn_ct_mul_scalar": (C.c_int, [vp, vp, C.c_uint32, C.c_uint64, vp, C.c_uint32]),
    "lmn_ct_mul_pt": (C.c_int, [vp, vp, C.c_uint32, vp, vp, C.c_uint32]),
    "lmn_ct_inner_sum": (C.c_int, [vp, vp, C.c_uint32, C.c_uint32]),
    "lmn_ct_rescale": (C.c_int, [vp, vp, C.c_uint32, vp, C.c_uint32]),
    "lmn_ct_copy": (C.c_int, [vp, vp, C.c_uint32, vp, C.c_uint32]),
    "lmn_limb_ntt": (C.c_int, [vp, vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int]),
}

_lib = None


class LumenosError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("liblumenos_b200 error %d: %s" % (code, msg))
        self.code = code


def lib():
    """The loaded C-ABI library.  Raises (never falls back) when liblumenos_b200.so is absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise ImportError("liblumenos_b200.so is not built (run `python -m lumenos_b200.build`); "
                              "lumenos_b200 has no CPU fallback")
        L = C.CDLL(SO_PATH)
        for name, (res, args) in SIGNATURES.items():
            f = getattr(L, name)   # AttributeError if the library does not export a declared symbol
            f.restype, f.argtypes = res, args
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise LumenosError(rc, lib().lmn_last_error().decode())
    return rc
