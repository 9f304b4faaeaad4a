"""ctypes binding of include/vvdec.h (libvvdec_b200.so).

This is exactly the stub a maintainer of the reference would add: the reference is
pure Python/PyTorch and has no FFI of its own, so the boundary is a C ABI with plain pointers and
sizes.  There is NO fallback: if the CUDA library is missing the import fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VVDEC_LIB", os.path.join(_HERE, "libvvdec_b200.so"))   # override: A/B builds only

VV_MAX_STAGES = 8
VV_F32, VV_BF16 = 0, 1
VV_LAYOUT_BTD, VV_LAYOUT_BDT = 0, 1
VV_OK = 0
ABI_VERSION = 2


class VvConfig(C.Structure):
    _fields_ = [
        ("vae_dim", C.c_int32),
        ("n_filters", C.c_int32),
        ("n_stages", C.c_int32),
        ("ratios", C.c_int32 * VV_MAX_STAGES),
        ("depths", C.c_int32 * VV_MAX_STAGES),
        ("kernel_size", C.c_int32),
        ("eps", C.c_float),
        ("device", C.c_int32),
        ("gemm_path", C.c_int32),
    ]


class VvSamplerConfig(C.Structure):
    _fields_ = [("hidden", C.c_int32), ("layers", C.c_int32), ("ffn", C.c_int32), ("latent", C.c_int32), ("eps", C.c_float),
                ("train_steps", C.c_int32), ("device", C.c_int32)]


class VvProfileRecord(C.Structure):
    _fields_ = [("kind", C.c_int32), ("stage", C.c_int32), ("launches", C.c_int64), ("ms", C.c_double)]


KERNEL_KINDS = ("prep", "stem", "mixer", "ffn1", "ffn2", "halo", "upsample", "head", "state", "ffn_fused",
                "block_fused", "enc_stem", "downsample", "pad", "enc_head", "ffn_pair")

# name -> (restype, argtypes): every symbol include/vvdec.h declares
_P, _I32, _I64 = C.c_void_p, C.c_int32, C.c_int64
SIGNATURES = {
    "vv_abi_version": (_I32, []),
    "vv_status_string": (C.c_char_p, [_I32]),
    "vv_last_error": (C.c_char_p, []),
    "vv_decoder_create": (_I32, [C.POINTER(VvConfig), C.POINTER(_P)]),
    "vv_decoder_destroy": (_I32, [_P]),
    "vv_decoder_load_tensor": (_I32, [_P, C.c_char_p, _P, _I32, C.POINTER(_I64), _I32, _I32]),
    "vv_decoder_finalize": (_I32, [_P]),
    "vv_decoder_param_count": (_I64, [_P]),
    "vv_decoder_set_option": (_I32, [_P, C.c_char_p, _I64]),
    "vv_encoder_set_option": (_I32, [_P, C.c_char_p, _I64]),
    "vv_state_check": (_I32, [_P]),
    "vv_decoder_hop": (_I32, [_P]),
    "vv_decoder_halo_frames": (_I32, [_P]),
    "vv_decoder_workspace_bytes": (_I64, [_P, _I32, _I32]),
    "vv_decode": (_I32, [_P, _P, _I32, _I32, _I32, _I32, _P, _I32, _P, _I64, _P, _P, _P]),
    "vv_decoder_take_launch_count": (_I64, [_P]),
    "vv_decoder_set_profiling": (_I32, [_P, _I32]),
    "vv_decoder_read_profile": (_I32, [_P, _P, _I32]),
    "vv_state_create": (_I32, [_P, _I32, C.POINTER(_P)]),
    "vv_state_destroy": (_I32, [_P]),
    "vv_state_n_slots": (_I32, [_P]),
    "vv_state_bytes_per_slot": (_I64, [_P]),
    "vv_state_zero": (_I32, [_P, _P, _I32, _P]),
    "vv_state_export": (_I32, [_P, _I32, _P, _I64]),
    "vv_state_import": (_I32, [_P, _I32, _P, _I64]),
    "vv_kernel_mixer": (_I32, [_P, _I32, _I32, _P, _P, _P, _I32, _I32, _P]),
    "vv_kernel_ffn": (_I32, [_P, _I32, _I32, _P, _P, _P, _P, _I32, _I32, _P]),
    "vv_kernel_block": (_I32, [_P, _I32, _I32, _P, _P, _I32, _I32, _P]),
    "vv_kernel_upsample": (_I32, [_P, _I32, _P, _P, _I32, _I32, _P]),
    "vv_test_gelu": (_I32, [_P, _P, _I32, _P]),
    "vv_kernel_head": (_I32, [_P, _P, _P, _I32, _I32, _I32, _P]),
    "vv_pcm16": (_I32, [_P, _I32, _I32, _I64, _I64, _P, _P, _P]),
    "vv_sampler_create": (_I32, [C.POINTER(VvSamplerConfig), C.POINTER(_P)]),
    "vv_sampler_destroy": (_I32, [_P]),
    "vv_sampler_load_tensor": (_I32, [_P, C.c_char_p, _P, _I32, C.POINTER(_I64), _I32, _I32]),
    "vv_sampler_finalize": (_I32, [_P]),
    "vv_sampler_param_count": (_I64, [_P]),
    "vv_sampler_prepare": (_I32, [_P, _I32]),
    "vv_sampler_workspace_bytes": (_I64, [_P, _I32]),
    "vv_sampler_sample": (_I32, [_P, _P, _P, _I32, _I32, _P, C.c_float, _I32, _P, _P, _I64, _P]),
    "vv_sampler_head": (_I32, [_P, _P, C.c_float, _P, _I32, _I32, _P, _P, _I64, _P]),
    "vv_sampler_take_launch_count": (_I64, [_P]),
    "vv_pcm16_static": (_I32, [_P, _I32, _I32, _I64, _I64, _P, _P]),
    "vv_encoder_create": (_I32, [C.POINTER(VvConfig), C.POINTER(_P)]),
    "vv_encoder_destroy": (_I32, [_P]),
    "vv_encoder_load_tensor": (_I32, [_P, C.c_char_p, _P, _I32, C.POINTER(_I64), _I32, _I32]),
    "vv_encoder_finalize": (_I32, [_P]),
    "vv_encoder_param_count": (_I64, [_P]),
    "vv_encoder_hop": (_I32, [_P]),
    "vv_encoder_halo_frames": (_I32, [_P]),
    "vv_encoder_frames": (_I64, [_P, _I64]),
    "vv_encoder_workspace_bytes": (_I64, [_P, _I32, _I64]),
    "vv_encoder_state_create": (_I32, [_P, _I32, C.POINTER(_P)]),
    "vv_encode": (_I32, [_P, _P, _I32, _I32, _I64, _P, _P, _I64, _P, _P, _P]),
    "vv_kernel_enc_stem": (_I32, [_P, _P, _I32, _P, _I32, _I64, _P]),
    "vv_kernel_downsample": (_I32, [_P, _I32, _P, _P, _I32, _I32, _P]),
    "vv_kernel_enc_head": (_I32, [_P, _P, _P, _I32, _I32, _P]),
    "vv_encoder_take_launch_count": (_I64, [_P]),
    "vv_encoder_set_profiling": (_I32, [_P, _I32]),
    "vv_encoder_read_profile": (_I32, [_P, _P, _I32]),
}

_lib = None


class VvError(RuntimeError):
    def __init__(self, status: int, what: str, detail: str):
        super().__init__(f"{what}: {detail} (status {status})")
        self.status = status


def load():
    """dlopen the library and bind every declared symbol.  Raises (never falls back) when missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: the sm_100a CUDA library has not been built "
            "(python -c 'import __graft_entry__ as g; g.build()').  vvdec has no CPU or PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the .so is stale
        fn.restype = res
        fn.argtypes = args
    got = lib.vv_abi_version()
    if got != ABI_VERSION:
        raise ImportError(f"libvvdec_b200.so ABI {got} != binding ABI {ABI_VERSION}; rebuild")
    _lib = lib
    return lib


def check(status: int, what: str):
    if status != VV_OK:
        lib = load()
        detail = lib.vv_last_error().decode() or lib.vv_status_string(status).decode()
        err = VvError(status, what, detail)
        if status in (-1, -2, -3, -4):           # invalid arg / shape / dtype / unsupported
            raise ValueError(str(err))
        if status == -7:                         # VV_ERR_SLOT: the reference's dict cache would raise KeyError/IndexError
            raise IndexError(str(err))
        raise err
