"""ctypes view of include/hj_b200.h (the C ABI in csrc/libhj_b200.so).

This module only loads the shared library and declares signatures; it holds no
algorithm.  There is no CPU fallback: if the library is missing the import of
the package fails, and without a CUDA device hj_create returns HJ_ECUDA.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libhj_b200.so")

HJ_OK, HJ_EINVAL, HJ_ENOMEM, HJ_ECUDA, HJ_EOVERFLOW, HJ_ESTATE = 0, -1, -2, -3, -4, -5
HJ_ALGO_ATOMIC, HJ_ALGO_NOCC, HJ_ALGO_RADIX, HJ_ALGO_ADAPTIVE, HJ_ALGO_SORTMERGE = 0, 1, 2, 3, 4
HJ_PATH_NONE, HJ_PATH_GLOBAL, HJ_PATH_RADIX, HJ_PATH_SORTMERGE = 0, 1, 2, 3

# every symbol include/hj_b200.h declares (tests check the library exports them all)
ABI_SYMBOLS = [
    "hj_config_default", "hj_create", "hj_destroy", "hj_abi_version", "hj_device_count",
    "hj_strerror", "hj_last_error", "hj_reserve", "hj_build", "hj_probe", "hj_join",
    "hj_sample", "hj_checksums", "hj_radix_partition", "hj_exchange_partition",
    "hj_generate", "hj_device_alloc", "hj_device_free", "hj_host_alloc", "hj_host_free",
    "hj_copy", "hj_flush_l2", "hj_sync",
    "hj_radix_plan", "hj_ipc_export", "hj_ipc_open", "hj_ipc_close", "hj_peer_hist", "hj_peer_scatter",
    "hj_join_pass1", "hj_join_materialize", "hj_peer_scatter_async",
    "hj_peer_scatter_regions", "hj_join_regions", "hj_reserve_staging", "hj_alloc_count",
    "hj_xchg_create", "hj_xchg_connect", "hj_xchg_start", "hj_xchg_finish", "hj_join_noalloc",
]
# host-side mirror of the reference interface (csrc/ref_entry.hpp, csrc/datagen.hpp)
HOST_SYMBOLS = ["hj_ref_pro", "hj_ref_run", "hj_ref_set_radix", "hj_ref_set_device", "hj_ref_last_status",
                "hj_generate_data", "hj_mc_relation_pk", "hj_mc_relation_fk", "hj_mc_relation_nonunique",
                "hj_mc_relation_fk_from_pk", "hj_mc_relation_pk_lshuffle", "hj_mc_relation_zipf"]


class hj_config(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("algo", C.c_int32), ("device", C.c_int32),
        ("probe_length", C.c_uint32), ("scale_output", C.c_uint32), ("num_partitions", C.c_uint32),
        ("passes", C.c_uint32), ("transaction_size", C.c_uint32), ("enable_probe", C.c_uint32),
        ("sample_rounds", C.c_uint32), ("key_mask", C.c_uint64), ("adaptive_threshold", C.c_double),
        ("pass1_bits", C.c_uint64), ("tuple_bytes", C.c_uint64), ("reserved", C.c_uint64 * 2),
    ]


class hj_result(C.Structure):
    _fields_ = [
        ("matches", C.c_uint64), ("checksum", C.c_uint64), ("window_matches", C.c_uint64),
        ("input_sum", C.c_uint64), ("table_sum", C.c_uint64), ("overflow_sum", C.c_uint64),
        ("conflicts", C.c_uint64), ("cas_retries", C.c_uint64), ("probe_steps", C.c_uint64),
        ("sampled", C.c_uint64), ("sectors", C.c_uint64), ("contention", C.c_double),
        ("path_taken", C.c_int32), ("radix_bits", C.c_uint32), ("passes", C.c_uint32),
        ("kernel_launches", C.c_uint32),
        ("us_total", C.c_float), ("us_h2d", C.c_float), ("us_clear", C.c_float), ("us_sample", C.c_float),
        ("us_build", C.c_float), ("us_probe", C.c_float), ("us_hist", C.c_float), ("us_scan", C.c_float),
        ("us_scatter1", C.c_float), ("us_scatter2", C.c_float), ("us_join", C.c_float),
        ("us_reserved", C.c_float * 5),
        ("hbm_bytes", C.c_uint64), ("max_partition", C.c_uint64), ("zero_keys", C.c_uint64),
        ("reserved", C.c_uint64 * 5),
    ]

    def as_dict(self):
        out = {}
        for name, _ in self._fields_:
            if "reserved" in name:
                continue
            out[name] = getattr(self, name)
        return out


def load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `make -C htm_hashjoin_b200/csrc` or "
            "`python -c 'import __graft_entry__ as g; g.build()'`.  There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    u64p, vp = C.c_void_p, C.c_void_p  # raw addresses: host numpy or device pointers alike
    H = C.c_void_p
    sig = {
        "hj_config_default": (None, [C.POINTER(hj_config)]),
        "hj_create": (C.c_int, [C.POINTER(hj_config), C.POINTER(H)]),
        "hj_destroy": (None, [H]),
        "hj_abi_version": (C.c_int, []),
        "hj_device_count": (C.c_int, []),
        "hj_strerror": (C.c_char_p, [C.c_int]),
        "hj_last_error": (C.c_char_p, [H]),
        "hj_reserve": (C.c_int, [H, C.c_uint64, C.c_uint64]),
        "hj_reserve_staging": (C.c_int, [H, C.c_uint64, C.c_uint64]),
        "hj_alloc_count": (C.c_uint64, [H]),
        "hj_build": (C.c_int, [H, u64p, C.c_uint64, C.POINTER(hj_result)]),
        "hj_probe": (C.c_int, [H, u64p, C.c_uint64, C.POINTER(hj_result)]),
        "hj_join": (C.c_int, [H, u64p, C.c_uint64, u64p, C.c_uint64, C.POINTER(hj_result)]),
        "hj_sample": (C.c_int, [H, u64p, C.c_uint64, C.POINTER(hj_result)]),
        "hj_checksums": (C.c_int, [H, u64p, C.c_uint64, C.POINTER(hj_result)]),
        "hj_radix_partition": (C.c_int, [H, u64p, C.c_uint64, C.c_uint32, C.c_uint32, u64p, u64p]),
        "hj_exchange_partition": (C.c_int, [H, u64p, C.c_uint64, C.c_uint32, C.c_uint32, u64p, u64p]),
        "hj_radix_plan": (C.c_int, [H, C.c_uint64, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32),
                                    C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
        "hj_ipc_export": (C.c_int, [H, vp, vp]),
        "hj_ipc_open": (C.c_int, [H, vp, C.POINTER(vp)]),
        "hj_ipc_close": (C.c_int, [H, vp]),
        "hj_peer_hist": (C.c_int, [H, u64p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, u64p]),
        "hj_peer_scatter": (C.c_int, [H, u64p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32,
                                      C.POINTER(C.c_uint64), u64p]),
        "hj_peer_scatter_async": (C.c_int, [H, u64p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32,
                                            C.POINTER(C.c_uint64), u64p, C.c_uint32]),
        "hj_peer_scatter_regions": (C.c_int, [H, u64p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32,
                                              C.POINTER(C.c_uint64), u64p, C.c_uint32, C.c_uint64, C.c_uint64,
                                              C.POINTER(C.c_uint32)]),
        "hj_join_regions": (C.c_int, [H, u64p, C.c_uint64, u64p, C.c_uint64, u64p, C.c_uint64, u64p, C.c_uint64,
                                      C.c_uint32, C.POINTER(hj_result)]),
        "hj_join_pass1": (C.c_int, [H, u64p, C.c_uint64, u64p, u64p, C.c_uint64, u64p, C.POINTER(hj_result)]),
        "hj_join_materialize": (C.c_int, [H, u64p, C.c_uint64, u64p, C.c_uint64, u64p, C.c_uint64,
                                          C.POINTER(C.c_uint64), C.POINTER(hj_result)]),
        "hj_xchg_create": (C.c_int, [H, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32,
                                     C.POINTER(vp), C.POINTER(C.c_uint64)]),
        "hj_xchg_connect": (C.c_int, [H, C.POINTER(C.c_uint64)]),
        "hj_xchg_start": (C.c_int, [H, u64p, u64p, C.c_uint64]),
        "hj_xchg_finish": (C.c_int, [H, C.POINTER(hj_result)]),
        "hj_join_noalloc": (C.c_int, [C.c_int, vp, u64p, C.c_uint64, u64p, C.c_uint64, C.c_uint32, u64p, C.c_char_p,
                                      C.c_uint32]),
        "hj_generate": (C.c_int, [H, C.c_int, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, u64p]),
        "hj_device_alloc": (C.c_int, [H, C.c_uint64, C.POINTER(vp)]),
        "hj_device_free": (C.c_int, [H, vp]),
        "hj_host_alloc": (C.c_int, [C.c_uint64, C.POINTER(vp)]),
        "hj_host_free": (C.c_int, [vp]),
        "hj_copy": (C.c_int, [H, vp, vp, C.c_uint64]),
        "hj_flush_l2": (C.c_int, [H]),
        "hj_sync": (C.c_int, [H]),
        "hj_ref_run": (C.c_int, [C.c_char_p, u64p, C.c_uint64, u64p, C.c_uint64, C.c_uint32, C.c_uint32,
                                 C.c_uint32, C.c_uint32, C.c_uint32, C.c_char_p, C.c_uint32]),
        "hj_ref_pro": (C.c_int64, [u64p, C.c_uint32, u64p, C.c_uint32, C.c_int]),
        "hj_ref_set_radix": (None, [C.c_uint32, C.c_uint32]),
        "hj_ref_set_device": (None, [C.c_int]),
        "hj_ref_last_status": (C.c_int, []),
        "hj_generate_data": (C.c_int, [C.c_char_p, C.c_uint64, C.c_uint64, C.c_int, u64p]),
        "hj_mc_relation_pk": (None, [u64p, C.c_uint64, C.c_uint]),
        "hj_mc_relation_fk": (None, [u64p, C.c_uint64, C.c_uint64, C.c_uint]),
        "hj_mc_relation_nonunique": (None, [u64p, C.c_uint64, C.c_uint64, C.c_uint, C.c_uint32]),
        "hj_mc_relation_fk_from_pk": (None, [u64p, C.c_uint64, u64p, C.c_uint64, C.c_uint]),
        "hj_mc_relation_pk_lshuffle": (None, [u64p, C.c_uint64, C.c_int, C.c_uint]),
        "hj_mc_relation_zipf": (None, [u64p, C.c_uint64, C.c_uint64, C.c_double, C.c_uint]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib
