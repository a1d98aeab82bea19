"""ctypes bindings of libp1gpu.so (include/p1gpu.h).  Plumbing only."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libp1gpu.so")

c_vp = ctypes.c_void_p
c_sz = ctypes.c_size_t
c_i = ctypes.c_int
c_u64 = ctypes.c_uint64
c_i64 = ctypes.c_int64
c_i32 = ctypes.c_int32
c_d = ctypes.c_double
c_cp = ctypes.c_char_p

P1_GRAM_AUTO, P1_GRAM_IMAD, P1_GRAM_TCGEN05 = 0, 1, 2
# kernel classes of p1_ctx_profile_read (enum in include/p1gpu.h)
(P1_K_ENCODE, P1_K_LAYOUT, P1_K_DIAG, P1_K_GRAM_IMAD, P1_K_GRAM_LIMBS, P1_K_GRAM_TC, P1_K_TI, P1_K_PARTY_B_MASK,
 P1_K_PARTY_A_STEP, P1_K_PARTY_B_SHARE, P1_K_COT_RECV, P1_K_COT_SEND, P1_K_COT_FINISH, P1_K_MISC, P1_K_MSG_PACK, P1_K_MSG_UNPACK, P1_K_SHARE_BITS,
 P1_K_TEXT, P1_K_MASKING_FUSED, P1_K_COUNT) = range(20)
P1_MSG_BAD_HEADER, P1_MSG_BAD_COUNT, P1_MSG_BAD_VARINT = 1, 2, 4

# every symbol include/p1gpu.h declares, with its signature
SIGNATURES = {
    "p1_ctx_create": (c_i, [c_i, c_i, c_i, ctypes.POINTER(c_vp)]),
    "p1_ctx_destroy": (None, [c_vp]),
    "p1_last_error": (c_cp, []),
    "p1_ctx_sync": (c_i, [c_vp]),
    "p1_ctx_wait": (c_i, [c_vp, c_vp]),
    "p1_ctx_stream": (c_vp, [c_vp]),
    "p1_ctx_launches": (c_u64, [c_vp]),
    "p1_ctx_profile": (c_i, [c_vp, c_i]),
    "p1_ctx_profile_read": (c_i, [c_vp, c_i, ctypes.POINTER(c_d), ctypes.POINTER(c_u64)]),
    "p1_kernel_class_name": (c_cp, [c_i]),
    "p1_ctx_set_option": (c_i, [c_vp, c_cp, ctypes.c_long]),
    "p1_dev_alloc": (c_i, [c_vp, c_sz, ctypes.POINTER(c_vp)]),
    "p1_dev_free": (c_i, [c_vp, c_vp]),
    "p1_dev_zero": (c_i, [c_vp, c_vp, c_sz]),
    "p1_host_alloc": (c_i, [c_vp, c_sz, ctypes.POINTER(c_vp)]),
    "p1_host_free": (c_i, [c_vp, c_vp]),
    "p1_copy_h2d": (c_i, [c_vp, c_vp, c_vp, c_sz]),
    "p1_copy_d2h": (c_i, [c_vp, c_vp, c_vp, c_sz]),
    "p1_copy_d2h_async": (c_i, [c_vp, c_vp, c_vp, c_sz]),
    "p1_ctx_mark": (c_i, [c_vp, c_i]),
    "p1_ctx_mark_sync": (c_i, [c_vp, c_i]),
    "p1_encode": (c_i, [c_vp, c_vp, c_vp, c_sz, c_sz, c_d, c_i, ctypes.POINTER(c_vp)]),
    "p1_data_from_words": (c_i, [c_vp, c_vp, c_vp, c_sz, c_sz, ctypes.POINTER(c_vp)]),
    "p1_data_free": (None, [c_vp]),
    "p1_data_download": (c_i, [c_vp, c_vp, c_vp, c_vp]),
    "p1_data_rows": (c_sz, [c_vp]),
    "p1_data_cols": (c_sz, [c_vp]),
    "p1_data_colmajor_dev": (c_vp, [c_vp, ctypes.POINTER(c_sz)]),
    "p1_acc_len": (c_sz, [c_sz]),
    "p1_local_block": (c_i, [c_vp, c_vp, c_sz, c_sz, c_i, c_vp, c_i]),
    "p1_gram_digits": (c_i, [c_vp, c_vp, c_sz, c_sz, c_i, ctypes.POINTER(c_i)]),
    "p1_diag_sequential": (c_i, [c_vp, c_vp, c_sz, c_sz, c_i, c_vp]),
    "p1_ti_generate": (c_i, [c_vp, c_cp, c_u64, c_sz, c_sz, c_sz, c_sz, c_vp, c_vp, c_sz, c_vp, c_vp]),
    "p1_ti_generate_keys": (c_i, [c_vp, c_cp, c_cp, c_cp, c_u64, c_sz, c_sz, c_sz, c_sz, c_vp, c_vp, c_sz, c_vp, c_vp]),
    "p1_ti_values": (c_i, [c_vp, c_vp, c_vp, c_sz, c_vp, c_vp]),
    "p1_party_b_mask": (c_i, [c_vp, c_vp, c_vp, c_sz, c_vp, c_sz, c_vp, c_sz]),
    "p1_party_a_step": (c_i, [c_vp, c_vp, c_vp, c_vp, c_sz, c_vp, c_sz, c_vp, c_sz, c_vp, c_sz, c_vp]),
    "p1_party_b_share": (c_i, [c_vp, c_vp, c_vp, c_vp, c_sz, c_vp, c_sz, c_vp]),
    "p1_acc_sub_values": (c_i, [c_vp, c_vp, c_vp, c_vp, c_sz]),
    "p1_masking_fused": (c_i, [c_vp, c_cp, c_u64, c_sz, c_sz, c_sz, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_sz, c_vp, c_vp,
                               c_vp, c_vp, c_vp]),
    "p1_cot_sender_new": (c_i, [c_vp, c_cp, c_vp, ctypes.POINTER(c_vp)]),
    "p1_cot_receiver_new": (c_i, [c_vp, c_vp, c_vp, ctypes.POINTER(c_vp)]),
    "p1_cot_free": (None, [c_vp]),
    "p1_cot_u_bytes": (c_sz, [c_vp, c_sz]),
    "p1_cot_c_bytes": (c_sz, [c_vp, c_sz]),
    "p1_cot_receiver_extend": (c_i, [c_vp, c_vp, c_vp, c_i32, c_sz, c_sz, c_vp]),
    "p1_cot_sender_step": (c_i, [c_vp, c_vp, c_vp, c_i32, c_sz, c_sz, c_vp, c_vp, c_vp, c_i64]),
    "p1_cot_advance": (c_i, [c_vp, c_sz, c_sz]),
    "p1_cot_set_cursor": (c_i, [c_vp, c_u64, c_u64]),
    "p1_cot_receiver_finish": (c_i, [c_vp, c_vp, c_vp, c_vp, c_i64]),
    "p1_probe_tensor_i8": (c_i, [c_vp, c_i, c_d, ctypes.POINTER(c_d), ctypes.POINTER(c_d), ctypes.POINTER(c_u64)]),
    "p1_finalize": (c_i, [c_vp, c_vp, c_sz, c_sz, c_sz, c_i, c_i, c_d, c_vp, c_vp,
                          ctypes.POINTER(c_vp), ctypes.POINTER(c_vp)]),
    "p1_finalize_async": (c_i, [c_vp, c_vp, c_sz, c_sz, c_vp, c_i, c_i, c_d, c_vp, c_vp, c_vp]),
    "p1_pool_trim": (None, []),
    "p1_encode_again": (c_i, [c_vp, c_vp, c_vp, c_vp, c_d, c_i]),
    "p1_group_create": (c_i, [c_vp, c_i, c_i, c_i, ctypes.POINTER(c_vp)]),
    "p1_group_destroy": (None, [c_vp]),
    "p1_group_size": (c_i, [c_vp]),
    "p1_group_ctx": (c_vp, [c_vp, c_i]),
    "p1_group_rows": (None, [c_vp, c_sz, c_i, ctypes.POINTER(c_sz), ctypes.POINTER(c_sz)]),
    "p1_group_set_option": (c_i, [c_vp, c_cp, ctypes.c_long]),
    "p1_group_sync": (c_i, [c_vp]),
    "p1_group_launches": (c_u64, [c_vp]),
    "p1_group_encode": (c_i, [c_vp, c_vp, c_vp, c_sz, c_sz, c_d, c_i, ctypes.POINTER(c_vp)]),
    "p1_gdata_free": (None, [c_vp]),
    "p1_gdata_shard": (c_vp, [c_vp, c_i]),
    "p1_group_acc_reset": (c_i, [c_vp, c_sz]),
    "p1_group_acc_dev": (c_vp, [c_vp, c_i]),
    "p1_group_reduce": (c_i, [c_vp]),
    "p1_group_local_block": (c_i, [c_vp, c_vp, c_sz, c_sz, c_i, c_i]),
    "p1_gvec_alloc": (c_i, [c_vp, c_sz, c_sz, ctypes.POINTER(c_vp)]),
    "p1_gvec_free": (None, [c_vp]),
    "p1_gvec_dev": (c_vp, [c_vp, c_i, ctypes.POINTER(c_sz)]),
    "p1_gvec_gather": (c_i, [c_vp, c_vp, c_sz, c_vp, c_sz]),
    "p1_gvec_scatter": (c_i, [c_vp, c_vp, c_sz, c_sz, c_vp]),
    "p1_group_ti_generate": (c_i, [c_vp, c_cp, c_u64, c_sz, c_sz, c_vp, c_vp, c_vp, c_vp]),
    "p1_group_ti_generate_keys": (c_i, [c_vp, c_cp, c_cp, c_cp, c_u64, c_sz, c_sz, c_vp, c_vp, c_vp, c_vp]),
    "p1_group_party_b_mask": (c_i, [c_vp, c_vp, c_vp, c_sz, c_vp, c_vp]),
    "p1_group_party_a_step": (c_i, [c_vp, c_vp, c_vp, c_vp, c_sz, c_vp, c_vp, c_vp]),
    "p1_group_party_b_share": (c_i, [c_vp, c_vp, c_vp, c_vp, c_sz, c_vp]),
    "p1_group_acc_sub_values": (c_i, [c_vp, c_sz, c_vp, c_vp, c_sz]),
    "p1_group_cot_sender_new": (c_i, [c_vp, c_cp, c_vp, ctypes.POINTER(c_vp)]),
    "p1_group_cot_receiver_new": (c_i, [c_vp, c_vp, c_vp, ctypes.POINTER(c_vp)]),
    "p1_group_cot_free": (None, [c_vp]),
    "p1_group_cot_receiver_extend": (c_i, [c_vp, c_vp, c_vp, c_i32, c_vp]),
    "p1_group_cot_sender_step": (c_i, [c_vp, c_vp, c_vp, c_i32, c_vp, c_vp, c_i64]),
    "p1_group_cot_receiver_finish": (c_i, [c_vp, c_vp, c_vp, c_vp, c_i64]),
    "p1_group_finalize": (c_i, [c_vp, c_vp, c_vp, c_sz, c_sz, c_i, c_i, c_d, ctypes.POINTER(c_vp), ctypes.POINTER(c_vp)]),
    "p1_msg_max_bytes": (c_sz, [c_vp, c_sz]),
    "p1_msg_pack": (c_i, [c_vp, c_vp, c_sz, c_sz, c_sz, c_vp, c_vp, c_sz, c_vp]),
    "p1_msg_pack_static": (c_i, [c_vp, c_vp, c_sz, c_sz, c_sz, c_vp, c_vp, c_sz, c_vp]),
    "p1_msg_unpack": (c_i, [c_vp, c_vp, c_sz, c_vp, c_sz, c_sz, c_vp, c_sz, c_vp, c_vp]),
    "p1_share_bits": (c_i, [c_vp, c_vp, c_sz, c_vp]),
    "p1_parse_doubles": (c_i, [c_vp, c_vp, c_sz, c_vp, c_sz, c_vp, c_vp, c_vp, c_sz]),
}


def build(verbose=False):
    """Compile libp1gpu.so for sm_100a (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-j8", "-C", os.path.join(HERE, "csrc")]
    if not verbose:
        cmd.insert(1, "-s")
    subprocess.check_call(cmd)
    return LIB_PATH


class P1Error(RuntimeError):
    pass


class P1Lib:
    """Loaded libp1gpu.so.  Raises if the library is missing: there is no fallback."""

    def __init__(self, path=LIB_PATH):
        if not os.path.exists(path):
            raise P1Error("libp1gpu.so not found at %s: build it with linreg_mpc_b200.build() "
                          "(the product path has no CPU fallback)" % path)
        self.path = path
        self.lib = ctypes.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(self.lib, name)  # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        self._libc = ctypes.CDLL(None)
        self._libc.free.argtypes = [c_vp]

    def check(self, status):
        if status != 0:
            raise P1Error(self.lib.p1_last_error().decode())

    def free(self, ptr):
        self._libc.free(ptr)

    def __getattr__(self, name):
        return getattr(self.lib, name)


def ptr(a):
    """void* of a numpy array / torch tensor / int / None"""
    if a is None:
        return None
    if isinstance(a, int):
        return c_vp(a)
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(c_vp)
    if hasattr(a, "data_ptr"):
        return c_vp(a.data_ptr())
    raise TypeError(type(a))
