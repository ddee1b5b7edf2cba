"""ctypes loader for libppq_b200.so (the C ABI declared in include/ppq_b200.h).

There is no CPU fallback: if the shared library is missing this module raises, and every compute entry point of
the library itself fails with PPQ_ECUDA when no CUDA device is present.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
# PPQ_B200_LIB lets tools/stage_times.py A/B-test differently compiled builds of the same library (still sm_100a CUDA).
LIB_PATH = os.environ.get("PPQ_B200_LIB") or os.path.join(_HERE, "lib", "libppq_b200.so")
CSRC = os.path.join(_HERE, "csrc")
HEADER = os.path.join(_HERE, "..", "include", "ppq_b200.h")

u64p = C.POINTER(C.c_uint64)
i64p = C.POINTER(C.c_int64)
i32p = C.POINTER(C.c_int32)
f64p = C.POINTER(C.c_double)
vp = C.c_void_p


class CtxInfo(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("logN", "L", "K", "dnum", "alpha", "device")]


def build(force=False):
    """Compile the CUDA extension in tree for sm_100a (nvcc cross-compiles without a GPU)."""
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".cpp", ".hpp", "Makefile"))]
    srcs.append(HEADER)
    stale = (not os.path.exists(LIB_PATH)) or any(os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", CSRC, "-j4"], stdout=subprocess.DEVNULL)
    return LIB_PATH


_SIGS = {
    # name: (restype, argtypes)
    "ppq_last_error": (C.c_char_p, []),
    "ppq_version": (C.c_char_p, []),
    "ppq_ctx_create_from_ccjson": (C.c_int, [C.c_char_p, C.c_size_t, C.c_int, C.POINTER(vp)]),
    "ppq_ctx_create": (C.c_int, [C.c_int, C.c_int, u64p, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]),
    "ppq_ctx_create_synthetic": (C.c_int, [C.c_int] * 7 + [C.POINTER(vp)]),
    "ppq_ctx_destroy": (None, [vp]),
    "ppq_ctx_info": (C.c_int, [vp, C.POINTER(CtxInfo)]),
    "ppq_ctx_moduli": (C.c_int, [vp, u64p]),
    "ppq_ctx_roots": (C.c_int, [vp, u64p]),
    "ppq_ctx_scaling_factor": (C.c_double, [vp, C.c_int]),
    "ppq_ctx_scaling_factor_big": (C.c_double, [vp, C.c_int]),
    "ppq_ctx_num_parts": (C.c_int, [vp, C.c_int]),
    "ppq_sync": (C.c_int, [vp, vp]),
    "ppq_ctx_set_wave": (C.c_int, [vp, C.c_int]),
    "ppq_ntt_forward": (C.c_int, [vp, vp, i32p, C.c_size_t, vp]),
    "ppq_ntt_inverse": (C.c_int, [vp, vp, i32p, C.c_size_t, vp]),
    "ppq_rekey_upload": (C.c_int, [vp, u64p, u64p, C.POINTER(vp)]),
    "ppq_rekey_from_device": (C.c_int, [vp, vp, vp, C.POINTER(vp)]),
    "ppq_rekey_destroy": (None, [vp]),
    "ppq_reencrypt": (C.c_int, [vp, vp, C.c_int, vp, vp, C.c_size_t, vp]),
    "ppq_reencrypt_accumulate": (C.c_int, [vp, vp, C.c_int, vp, vp, C.c_size_t, vp]),
    "ppq_ks_modup": (C.c_int, [vp, C.c_int, vp, vp, C.c_size_t, vp]),
    "ppq_ks_inner_product": (C.c_int, [vp, vp, C.c_int, vp, vp, C.c_size_t, vp]),
    "ppq_ks_moddown": (C.c_int, [vp, C.c_int, vp, vp, C.c_size_t, vp]),
    "ppq_params_dump": (C.c_int, [C.c_int, C.c_int, u64p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_size_t,
                                  C.POINTER(C.c_size_t)]),
    "ppq_params_chain": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, u64p, C.c_int, C.POINTER(C.c_int)]),
    "ppq_eval_add": (C.c_int, [vp, C.c_int, vp, vp, vp, C.c_size_t, vp]),
    "ppq_finalize_sum": (C.c_int, [vp, C.c_int, vp, C.c_size_t, vp]),
    "ppq_eval_mult_const": (C.c_int, [vp, C.c_int, vp, vp, C.c_size_t, C.c_double, C.c_double, C.c_int, vp]),
    "ppq_encode": (C.c_int, [vp, C.c_int, vp, C.c_size_t, C.c_double, vp, C.c_size_t, vp]),
    "ppq_encrypt": (C.c_int, [vp, C.c_int, vp, vp, vp, vp, vp, C.c_size_t, vp]),
    "ppq_encrypt_noalloc": (C.c_int, [vp, C.c_int, vp, vp, vp, vp, vp, C.c_size_t, vp]),
    "ppq_encode_encrypt": (C.c_int, [vp, C.c_int, vp, vp, vp, C.c_size_t, C.c_double, vp, vp, C.c_size_t, vp]),
    "ppq_sample_encrypt_noise": (C.c_int, [vp, C.c_char_p, vp, C.c_size_t, vp]),
    "ppq_chacha20_block": (C.c_int, [C.POINTER(C.c_uint32), C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "ppq_reencrypt_fanout": (C.c_int, [vp, C.POINTER(vp), C.c_int, C.c_int, vp, C.POINTER(vp), C.c_size_t, vp]),
    "ppq_keygen": (C.c_int, [vp, vp, vp, vp, vp, vp, vp]),
    "ppq_sample_keygen": (C.c_int, [vp, C.c_char_p, vp, vp, vp, vp]),
    "ppq_rekeygen": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, vp]),
    "ppq_decrypt_core": (C.c_int, [vp, C.c_int, vp, vp, vp, C.c_size_t, vp]),
    "ppq_decode": (C.c_int, [vp, C.c_int, vp, C.c_double, vp, C.c_size_t, vp]),
    "ppq_reencrypt_aggregate_host": (C.c_int, [vp, C.POINTER(vp), C.c_int, C.POINTER(vp), C.c_size_t, C.c_size_t, vp]),
    "ppq_agg_create": (C.c_int, [vp, C.c_int, C.c_size_t, C.POINTER(vp)]),
    "ppq_agg_add": (C.c_int, [vp, vp, vp, vp]),
    "ppq_agg_add_many": (C.c_int, [vp, C.POINTER(vp), C.POINTER(vp), C.c_size_t, vp]),
    "ppq_agg_finish": (C.c_int, [vp, vp, vp]),
    "ppq_agg_count": (C.c_size_t, [vp]),
    "ppq_agg_destroy": (None, [vp]),
    "ppq_comm_unique_id": (C.c_int, [C.c_char_p]),
    "ppq_comm_init": (C.c_int, [vp, C.c_int, C.c_int, C.c_char_p, C.POINTER(vp)]),
    "ppq_comm_from_nccl": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]),
    "ppq_comm_destroy": (None, [vp]),
    "ppq_reduce_across_ranks": (C.c_int, [vp, vp, C.c_int, vp, C.c_size_t, C.c_int, vp]),
    "ppq_reencrypt_aggregate_host_ranks": (C.c_int, [vp, vp, C.c_int, C.POINTER(vp), C.c_int, C.POINTER(vp), C.c_size_t,
                                                     C.c_size_t, vp]),
    "ppq_copy_ceiling_host": (C.c_int, [vp, C.c_int, C.POINTER(vp), C.c_size_t, C.c_size_t, vp]),
    "ppq_packed_words": (C.c_size_t, [vp, C.c_int]),
    "ppq_pack_host": (C.c_int, [vp, C.c_int, vp, vp, C.c_size_t]),
    "ppq_unpack_host": (C.c_int, [vp, C.c_int, vp, vp, C.c_size_t]),
    "ppq_pack": (C.c_int, [vp, C.c_int, vp, vp, C.c_size_t, vp]),
    "ppq_unpack": (C.c_int, [vp, C.c_int, vp, vp, C.c_size_t, vp]),
    "ppq_reencrypt_aggregate_host_packed": (C.c_int, [vp, C.POINTER(vp), C.c_int, C.POINTER(vp), C.c_size_t, C.c_size_t,
                                                      vp]),
    "ppq_reencrypt_aggregate_host_packed_ranks": (C.c_int, [vp, vp, C.c_int, C.POINTER(vp), C.c_int, C.POINTER(vp),
                                                            C.c_size_t, C.c_size_t, vp]),
    "ppq_copy_ceiling_host_packed": (C.c_int, [vp, C.c_int, C.POINTER(vp), C.c_size_t, C.c_size_t, vp]),
    "ppq_plan_shards": (C.c_int, [C.c_size_t, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_size_t)]),
    "ppq_launch_count": (C.c_uint64, []),
    "ppq_profile_enable": (C.c_int, [vp, C.c_int]),
    "ppq_profile_read": (C.c_int, [vp, f64p, u64p, C.c_int]),
    "ppq_profile_stage_name": (C.c_char_p, [C.c_int]),
}

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(ppqsflhe_b200 has no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


class PPQError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        raise PPQError(f"libppq_b200 error {rc}: {lib().ppq_last_error().decode()}")
