"""ctypes binding of include/iree_b200_hal.h and include/iree_b200_ops.h.

Host-side mirror of the reference's HAL objects for the dispatch path (device, executable
cache, executable, command buffer, semaphore, channel), thin enough that every method is
one C-ABI call. Errors surface as `Status` exceptions carrying the reference's status code
names. Loading the library never needs a GPU; creating a `Device` does, and fails loudly
(`Status: UNAVAILABLE`) without one — there is no CPU fallback anywhere in this package.
"""

from __future__ import annotations

import ctypes
import os
from ctypes import (POINTER, Structure, byref, c_char_p, c_int32, c_int64, c_size_t, c_uint32,
                    c_uint64, c_void_p)

from . import build as _build

UINT64_MAX = (1 << 64) - 1

I8, F16, BF16, F32, F8E4M3FN, F8E5M2 = 0, 1, 2, 3, 4, 5
ELEMENT_TYPES = {"i8": I8, "f16": F16, "bf16": BF16, "f32": F32, "f8e4m3fn": F8E4M3FN,
                 "f8e5m2": F8E5M2}

MATMUL_ACCUMULATE = 1 << 0
MATMUL_TRANSPOSE_RHS = 1 << 1
MATMUL_TRUNCATE_RESULT = 1 << 2
MATMUL_BIAS = 1 << 3
MATMUL_RELU = 1 << 4
MATMUL_FORCE_GENERIC = 1 << 8
MATMUL_FORCE_1CTA = 1 << 9
MATMUL_FORCE_2CTA = 1 << 10
MATMUL_MX_EXPAND = 1 << 11
MATMUL_MMT4D_RELAYOUT = 1 << 12
MATMUL_DEVICE_MAPS = 1 << 13

ATTENTION_MASK_NONE, ATTENTION_MASK_BOOL, ATTENTION_MASK_F32, ATTENTION_MASK_CAUSAL = 0, 1, 2, 3
ATTENTION_FORCE_GENERIC = 1 << 8
ATTENTION_FORCE_PREFILL = 1 << 9
ATTENTION_FORCE_DECODE = 1 << 10
ATTENTION_FORCE_DECODE_SPLIT = 1 << 11
ATTENTION_FORCE_DECODE_CLUSTER = 1 << 12
ATTENTION_EXP2_MUFU_ONLY = 1 << 13
ATTENTION_EXP2_POLY_25 = 1 << 14
ATTENTION_EXP2_POLY_50 = 1 << 15
ATTENTION_NO_TAIL_SPLIT = 1 << 16
ATTENTION_DEVICE_MAPS = 1 << 17

MODE_GRAPH, MODE_STREAM = 0, 1
MEMORY_DEVICE_LOCAL, MEMORY_HOST_LOCAL = 1, 2

# iree_hal_collective_* encodings (hal/command_buffer.h:300-462).
ALL_GATHER, ALL_REDUCE, ALL_TO_ALL, BROADCAST, REDUCE, REDUCE_SCATTER, SEND, RECV, SEND_RECV = range(9)
REDUCTION_NONE, REDUCTION_SUM, REDUCTION_PRODUCT, REDUCTION_MIN, REDUCTION_MAX, REDUCTION_AVG = range(6)
COLLECTIVE_ELEMENT_TYPES = {"i8": 0, "u8": 1, "i16": 2, "u16": 3, "i32": 4, "u32": 5, "i64": 6,
                            "u64": 7, "f16": 8, "f32": 9, "f64": 10, "bf16": 11}
CHANNEL_ID_SIZE = 128


def collective_op(kind: int, reduction: int, element_type) -> int:
    if isinstance(element_type, str):
        element_type = COLLECTIVE_ELEMENT_TYPES[element_type]
    return kind | (reduction << 8) | (element_type << 16)


class Status(RuntimeError):
    """A non-OK iree_b200_status_t (codes are iree_status_code_t values)."""

    def __init__(self, code: int, name: str, message: str):
        super().__init__(f"{name}: {message}")
        self.code = code
        self.name = name
        self.message = message


class DeviceParams(Structure):
    _fields_ = [("command_buffer_mode", c_int32), ("reserved0", c_int32),
                ("external_dispatch_stream", c_uint64), ("overlap_collectives", c_int32),
                ("reserved1", c_int32)]


class BufferRef(Structure):
    _fields_ = [("buffer", c_void_p), ("offset", c_uint64), ("length", c_uint64)]


class ExportInfo(Structure):
    _fields_ = [("kernel_name", c_char_p), ("block_dims", c_uint32 * 3),
                ("block_shared_memory_size", c_uint32), ("constant_count", c_uint32),
                ("binding_count", c_uint32), ("tma_map_count", c_uint32), ("reserved", c_uint32)]


class ChannelParams(Structure):
    _fields_ = [("flags", c_uint64), ("id", c_void_p), ("id_length", c_size_t),
                ("rank", c_int32), ("count", c_int32)]


class SymmetricRegionParams(Structure):
    _fields_ = [("byte_length", c_uint64), ("peer_bases", POINTER(c_uint64)),
                ("multicast_base", c_uint64), ("peer_flag_blocks", POINTER(c_uint64)),
                ("max_blocks", c_int32)]


SYMMETRIC_FLAG_BYTES = 2048


class MatmulSelection(Structure):
    _fields_ = [("export_ordinal", c_int32), ("workgroup_count", c_uint32 * 3),
                ("constants", c_uint32 * 5)]


class AttentionDispatch(Structure):
    _fields_ = [("export_ordinal", c_int32), ("workgroup_count", c_uint32 * 3),
                ("constant_count", c_uint32), ("constants", c_uint32 * 8)]


class AttentionSelection(Structure):
    _fields_ = [("dispatch_count", c_int32), ("dispatches", AttentionDispatch * 2),
                ("workspace_bytes", c_uint64)]


_STATUS_FUNCS = {
    # name: argtypes (all return iree_b200_status_t)
    "iree_b200_device_create": [c_int32, POINTER(DeviceParams), POINTER(c_void_p)],
    "iree_b200_device_query_i64": [c_void_p, c_char_p, c_char_p, POINTER(c_int64)],
    "iree_b200_device_wait_idle": [c_void_p],
    "iree_b200_allocator_allocate_buffer": [c_void_p, c_int32, c_uint64, POINTER(c_void_p)],
    "iree_b200_allocator_import_buffer": [c_void_p, c_uint64, c_uint64, POINTER(c_void_p)],
    "iree_b200_device_transfer_h2d": [c_void_p, c_void_p, c_void_p, c_uint64, c_uint64],
    "iree_b200_device_transfer_d2h": [c_void_p, c_void_p, c_uint64, c_void_p, c_uint64],
    "iree_b200_executable_cache_create": [c_void_p, c_char_p, POINTER(c_void_p)],
    "iree_b200_executable_infer_format": [c_void_p, c_size_t, c_size_t, c_char_p,
                                          POINTER(c_size_t)],
    "iree_b200_executable_cache_prepare_executable": [c_void_p, c_char_p, c_void_p, c_size_t,
                                                      POINTER(c_void_p)],
    "iree_b200_executable_export_info": [c_void_p, c_int32, POINTER(ExportInfo)],
    "iree_b200_executable_lookup_export_by_name": [c_void_p, c_char_p, POINTER(c_int32)],
    "iree_b200_command_buffer_create": [c_void_p, POINTER(c_void_p)],
    "iree_b200_command_buffer_begin": [c_void_p],
    "iree_b200_command_buffer_end": [c_void_p],
    "iree_b200_command_buffer_execution_barrier": [c_void_p],
    "iree_b200_command_buffer_fill_buffer": [c_void_p, BufferRef, c_void_p, c_size_t],
    "iree_b200_command_buffer_update_buffer": [c_void_p, c_void_p, c_size_t, BufferRef],
    "iree_b200_command_buffer_copy_buffer": [c_void_p, BufferRef, BufferRef],
    "iree_b200_command_buffer_dispatch": [c_void_p, c_void_p, c_int32, POINTER(c_uint32),
                                          POINTER(c_uint32), c_size_t, POINTER(BufferRef),
                                          c_size_t, c_uint64],
    "iree_b200_command_buffer_collective": [c_void_p, c_void_p, c_uint32, c_uint32, BufferRef,
                                            BufferRef, c_uint64],
    "iree_b200_semaphore_create": [c_void_p, c_uint64, POINTER(c_void_p)],
    "iree_b200_semaphore_query": [c_void_p, POINTER(c_uint64)],
    "iree_b200_semaphore_wait": [c_void_p, c_uint64, c_int64],
    "iree_b200_semaphore_signal": [c_void_p, c_uint64],
    "iree_b200_semaphore_fail": [c_void_p, ctypes.c_int, ctypes.c_char_p],
    "iree_b200_device_queue_execute": [c_void_p, c_size_t, POINTER(c_void_p), POINTER(c_uint64),
                                       c_size_t, POINTER(c_void_p), POINTER(c_uint64), c_void_p],
    "iree_b200_device_queue_alloca": [c_void_p, c_size_t, POINTER(c_void_p), POINTER(c_uint64),
                                      c_size_t, POINTER(c_void_p), POINTER(c_uint64), c_int32,
                                      c_uint64, POINTER(c_void_p)],
    "iree_b200_device_queue_dealloca": [c_void_p, c_size_t, POINTER(c_void_p), POINTER(c_uint64),
                                        c_size_t, POINTER(c_void_p), POINTER(c_uint64), c_void_p],
    "iree_b200_channel_id_create": [c_void_p],
    "iree_b200_channel_provider_query_default_rank_and_count": [POINTER(c_int32),
                                                                POINTER(c_int32)],
    "iree_b200_channel_provider_exchange_default_id": [c_int32, c_void_p],
    "iree_b200_channel_query_collective_status": [c_void_p, POINTER(ctypes.c_uint32)],
    "iree_b200_channel_create": [c_void_p, POINTER(ChannelParams), POINTER(c_void_p)],
    "iree_b200_channel_split": [c_void_p, c_int32, c_int32, c_uint64, POINTER(c_void_p)],
    "iree_b200_channel_attach_symmetric_region": [c_void_p, c_void_p,
                                                  POINTER(SymmetricRegionParams)],
    "iree_b200_ops_create": [c_void_p, POINTER(c_void_p)],
    "iree_b200_ops_create_from_data": [c_void_p, c_void_p, c_size_t, POINTER(c_void_p)],
    "iree_b200_ops_create_offline": [c_void_p, c_size_t, c_int32, POINTER(c_void_p)],
    "iree_b200_ops_set_reserved_sms": [c_void_p, c_int32],
    "iree_b200_ops_set_host_pipeline": [c_void_p, c_uint64, c_int32, c_int32],
    "iree_b200_ops_select_matmul": [c_void_p, c_int32, c_int64, c_int64, c_int64, c_int64,
                                    c_uint32, c_uint32, POINTER(MatmulSelection)],
    "iree_b200_ops_record_matmul": [c_void_p, c_void_p, c_int32, c_int64, c_int64, c_int64,
                                    c_int64, c_uint32, BufferRef, BufferRef, BufferRef],
    "iree_b200_ops_matmul_host": [c_void_p, c_int32, c_int64, c_int64, c_int64, c_int64,
                                  c_uint32, c_void_p, c_void_p, c_void_p],
    "iree_b200_ops_record_matmul_bias": [c_void_p, c_void_p, c_int32, c_int64, c_int64, c_int64,
                                         c_int64, c_uint32, BufferRef, BufferRef, BufferRef,
                                         BufferRef],
    "iree_b200_ops_matmul_bias_host": [c_void_p, c_int32, c_int64, c_int64, c_int64, c_int64,
                                       c_uint32, c_void_p, c_void_p, c_void_p, c_void_p],
    "iree_b200_ops_record_matmul_mxfp4": [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_uint32,
                                          BufferRef, BufferRef, BufferRef, BufferRef, BufferRef],
    "iree_b200_ops_record_matmul_mxfp4_tensor": [c_void_p, c_void_p, c_int64, c_int64, c_int64,
                                                 c_uint32, BufferRef, BufferRef, BufferRef,
                                                 BufferRef, BufferRef, BufferRef],
    "iree_b200_ops_matmul_mxfp4_host": [c_void_p, c_int64, c_int64, c_int64, c_uint32, c_void_p,
                                        c_void_p, c_void_p, c_void_p, c_void_p],
    "iree_b200_ops_record_mmt4d": [c_void_p, c_void_p, c_int32, c_int64, c_int64, c_int64, c_int32,
                                   c_int32, c_int32, c_uint32, BufferRef, BufferRef, BufferRef],
    "iree_b200_ops_record_mmt4d_tensor": [c_void_p, c_void_p, c_int32, c_int64, c_int64, c_int64,
                                          c_int32, c_int32, c_int32, c_uint32, BufferRef,
                                          BufferRef, BufferRef, BufferRef],
    "iree_b200_ops_mmt4d_host": [c_void_p, c_int32, c_int64, c_int64, c_int64, c_int32, c_int32,
                                 c_int32, c_uint32, c_void_p, c_void_p, c_void_p],
    "iree_b200_ops_select_attention": [c_void_p, c_int32, c_int64, c_int64, c_int64, c_int64,
                                       c_int64, ctypes.c_float, c_int32, c_uint32, c_uint32,
                                       POINTER(AttentionSelection)],
    "iree_b200_ops_record_attention": [c_void_p, c_void_p, c_int32, c_int64, c_int64, c_int64,
                                       c_int64, c_int64, ctypes.c_float, c_int32, c_uint32,
                                       BufferRef, BufferRef, BufferRef, BufferRef, BufferRef,
                                       BufferRef],
    "iree_b200_ops_attention_host": [c_void_p, c_int32, c_int64, c_int64, c_int64, c_int64,
                                     c_int64, ctypes.c_float, c_int32, c_uint32, c_void_p,
                                     c_void_p, c_void_p, c_void_p, c_void_p],
    "iree_b200_attention_decode_no_alloc": [c_int32, c_int32, c_int64, c_int64, c_int64,
                                            ctypes.c_float, c_void_p, c_void_p, c_void_p, c_void_p,
                                            c_void_p],
}

_OTHER_FUNCS = {
    # name: (restype, argtypes)
    "iree_b200_status_message": (c_char_p, []),
    "iree_b200_status_code_string": (c_char_p, [c_int32]),
    "iree_b200_device_params_initialize": (None, [POINTER(DeviceParams)]),
    "iree_b200_device_retain": (None, [c_void_p]),
    "iree_b200_device_release": (None, [c_void_p]),
    "iree_b200_device_dispatch_stream": (c_uint64, [c_void_p]),
    "iree_b200_buffer_retain": (None, [c_void_p]),
    "iree_b200_buffer_release": (None, [c_void_p]),
    "iree_b200_buffer_device_pointer": (c_uint64, [c_void_p]),
    "iree_b200_buffer_host_pointer": (c_void_p, [c_void_p]),
    "iree_b200_buffer_byte_length": (c_uint64, [c_void_p]),
    "iree_b200_executable_cache_release": (None, [c_void_p]),
    "iree_b200_executable_cache_can_prepare_format": (c_int32, [c_void_p, c_char_p]),
    "iree_b200_executable_retain": (None, [c_void_p]),
    "iree_b200_executable_release": (None, [c_void_p]),
    "iree_b200_executable_export_count": (c_int32, [c_void_p]),
    "iree_b200_command_buffer_retain": (None, [c_void_p]),
    "iree_b200_command_buffer_release": (None, [c_void_p]),
    "iree_b200_semaphore_release": (None, [c_void_p]),
    "iree_b200_channel_retain": (None, [c_void_p]),
    "iree_b200_channel_release": (None, [c_void_p]),
    "iree_b200_channel_query_collective_counts": (None, [c_void_p, POINTER(c_uint64),
                                                         POINTER(c_uint64)]),
    "iree_b200_channel_query_rank_and_count": (None, [c_void_p, POINTER(c_int32),
                                                      POINTER(c_int32)]),
    "iree_b200_ops_builtin_executable_data": (c_void_p, [POINTER(c_size_t)]),
    "iree_b200_ops_release": (None, [c_void_p]),
    "iree_b200_ops_matmul_mxfp4_workspace_size": (c_uint64, [c_int64, c_int64, c_int64]),
    "iree_b200_ops_matmul_mxfp4_scale_workspace_size": (c_uint64, [c_int64, c_int64, c_int64]),
    "iree_b200_ops_mmt4d_workspace_size": (c_uint64, [c_int32, c_int64, c_int64, c_int64, c_int32,
                                                      c_int32, c_int32]),
    "iree_b200_ops_executable": (c_void_p, [c_void_p]),
    "iree_b200_ops_dispatch_count": (c_uint64, [c_void_p]),
}

ALL_SYMBOLS = sorted(list(_STATUS_FUNCS) + list(_OTHER_FUNCS))

_lib = None


def lib() -> ctypes.CDLL:
    """Loads iree_b200/_build/libiree_b200.so (building it first if sources changed)."""
    global _lib
    if _lib is None:
        path = _build.lib_path()
        if not os.path.exists(path) or os.environ.get("IREE_B200_REBUILD"):
            path = _build.build()
        l = ctypes.CDLL(path)
        for name, argtypes in _STATUS_FUNCS.items():
            fn = getattr(l, name)
            fn.restype = c_int32
            fn.argtypes = argtypes
        for name, (restype, argtypes) in _OTHER_FUNCS.items():
            fn = getattr(l, name)
            fn.restype = restype
            fn.argtypes = argtypes
        _lib = l
    return _lib


def check(status: int):
    if status != 0:
        l = lib()
        raise Status(status, l.iree_b200_status_code_string(status).decode(),
                     l.iree_b200_status_message().decode(errors="replace"))


def _ptr(obj):
    return obj._h if obj is not None else None


NO_ALLOC_SCRATCH_BYTES = 4 << 20   # IREE_B200_NO_ALLOC_SCRATCH_BYTES


def attention_decode_no_alloc(device_ordinal: int, element_type, B, M, K2, scale, q_ptr: int,
                              k_ptr: int, v_ptr: int, out_ptr: int, stream: int = 0):
    """iree_b200_attention_decode_no_alloc: split-KV decode attention on DEVICE pointers whose
    scratch is a static device array inside the library — no allocation anywhere in the call.
    Asynchronous on |stream| (a CUstream / cudaStream_t handle; 0 = the default stream)."""
    check(lib().iree_b200_attention_decode_no_alloc(
        device_ordinal, _etype(element_type), B, M, K2, scale, q_ptr or None, k_ptr or None,
        v_ptr or None, out_ptr or None, stream or None))


class Buffer:
    def __init__(self, handle, device):
        self._h = handle
        self.device = device

    @property
    def device_pointer(self) -> int:
        return lib().iree_b200_buffer_device_pointer(self._h)

    @property
    def byte_length(self) -> int:
        return lib().iree_b200_buffer_byte_length(self._h)

    def ref(self, offset=0, length=UINT64_MAX) -> BufferRef:
        return BufferRef(self._h, offset, length)

    def release(self):
        if self._h:
            lib().iree_b200_buffer_release(self._h)
            self._h = None

    __del__ = release


class Device:
    """iree_hal_device_t for one B200 (include/iree_b200_hal.h, Device section)."""

    def __init__(self, ordinal=0, mode=MODE_STREAM, external_stream=0, overlap_collectives=False):
        params = DeviceParams()
        lib().iree_b200_device_params_initialize(byref(params))
        params.command_buffer_mode = mode
        params.external_dispatch_stream = external_stream
        params.overlap_collectives = int(overlap_collectives)
        h = c_void_p()
        check(lib().iree_b200_device_create(ordinal, byref(params), byref(h)))
        self._h = h
        self.ordinal = ordinal

    def query_i64(self, category: str, key: str) -> int:
        v = c_int64()
        check(lib().iree_b200_device_query_i64(self._h, category.encode(), key.encode(), byref(v)))
        return v.value

    @property
    def stream(self) -> int:
        return lib().iree_b200_device_dispatch_stream(self._h)

    def wait_idle(self):
        check(lib().iree_b200_device_wait_idle(self._h))

    def allocate(self, nbytes: int, memory_type=MEMORY_DEVICE_LOCAL) -> Buffer:
        h = c_void_p()
        check(lib().iree_b200_allocator_allocate_buffer(self._h, memory_type, nbytes, byref(h)))
        return Buffer(h, self)

    def import_buffer(self, device_pointer: int, nbytes: int) -> Buffer:
        h = c_void_p()
        check(lib().iree_b200_allocator_import_buffer(self._h, device_pointer, nbytes, byref(h)))
        return Buffer(h, self)

    def h2d(self, host_ptr: int, target: Buffer, nbytes: int, offset=0):
        check(lib().iree_b200_device_transfer_h2d(self._h, host_ptr, target._h, offset, nbytes))

    def d2h(self, source: Buffer, host_ptr: int, nbytes: int, offset=0):
        check(lib().iree_b200_device_transfer_d2h(self._h, source._h, offset, host_ptr, nbytes))

    @staticmethod
    def _semaphore_lists(waits, signals):
        nw, ns = len(waits), len(signals)
        ws = (c_void_p * max(nw, 1))(*[s._h for s, _ in waits])
        wv = (c_uint64 * max(nw, 1))(*[v for _, v in waits])
        ss = (c_void_p * max(ns, 1))(*[s._h for s, _ in signals])
        sv = (c_uint64 * max(ns, 1))(*[v for _, v in signals])
        return nw, ws, wv, ns, ss, sv

    def execute(self, cb, waits=(), signals=()):
        check(lib().iree_b200_device_queue_execute(self._h, *self._semaphore_lists(waits, signals),
                                                   _ptr(cb)))

    def queue_alloca(self, nbytes: int, waits=(), signals=(),
                     memory_type=MEMORY_DEVICE_LOCAL) -> "Buffer":
        h = c_void_p()
        check(lib().iree_b200_device_queue_alloca(self._h, *self._semaphore_lists(waits, signals),
                                                  memory_type, nbytes, byref(h)))
        return Buffer(h, self)

    def queue_dealloca(self, buffer: "Buffer", waits=(), signals=()):
        check(lib().iree_b200_device_queue_dealloca(
            self._h, *self._semaphore_lists(waits, signals), buffer._h))

    def release(self):
        if getattr(self, "_h", None):
            lib().iree_b200_device_release(self._h)
            self._h = None

    __del__ = release


class Semaphore:
    def __init__(self, device: Device, initial=0):
        h = c_void_p()
        check(lib().iree_b200_semaphore_create(device._h, initial, byref(h)))
        self._h = h

    def query(self) -> int:
        v = c_uint64()
        check(lib().iree_b200_semaphore_query(self._h, byref(v)))
        return v.value

    def wait(self, value: int, timeout_ns=-1):
        check(lib().iree_b200_semaphore_wait(self._h, value, timeout_ns))

    def signal(self, value: int):
        check(lib().iree_b200_semaphore_signal(self._h, value))

    def fail(self, code: int, message: str = None):
        check(lib().iree_b200_semaphore_fail(self._h, code,
                                             None if message is None else message.encode()))

    def release(self):
        if getattr(self, "_h", None):
            lib().iree_b200_semaphore_release(self._h)
            self._h = None

    __del__ = release


class Executable:
    def __init__(self, handle, owned=True):
        self._h = handle
        self._owned = owned

    @property
    def export_count(self) -> int:
        return lib().iree_b200_executable_export_count(self._h)

    def export_info(self, ordinal: int) -> dict:
        info = ExportInfo()
        check(lib().iree_b200_executable_export_info(self._h, ordinal, byref(info)))
        return {"kernel_name": info.kernel_name.decode(), "block_dims": list(info.block_dims),
                "block_shared_memory_size": info.block_shared_memory_size,
                "constant_count": info.constant_count, "binding_count": info.binding_count,
                "tma_map_count": info.tma_map_count}

    def lookup(self, name: str) -> int:
        o = c_int32()
        check(lib().iree_b200_executable_lookup_export_by_name(self._h, name.encode(), byref(o)))
        return o.value

    def release(self):
        if getattr(self, "_h", None) and self._owned:
            lib().iree_b200_executable_release(self._h)
        self._h = None

    __del__ = release


class ExecutableCache:
    def __init__(self, device: Device, identifier="default"):
        h = c_void_p()
        check(lib().iree_b200_executable_cache_create(device._h, identifier.encode(), byref(h)))
        self._h = h

    def can_prepare_format(self, fmt: str) -> bool:
        return bool(lib().iree_b200_executable_cache_can_prepare_format(self._h, fmt.encode()))

    def prepare(self, data: bytes, fmt="PTXE") -> Executable:
        h = c_void_p()
        buf = ctypes.create_string_buffer(data, len(data))
        check(lib().iree_b200_executable_cache_prepare_executable(self._h, fmt.encode(), buf,
                                                                 len(data), byref(h)))
        return Executable(h)

    def release(self):
        if getattr(self, "_h", None):
            lib().iree_b200_executable_cache_release(self._h)
            self._h = None

    __del__ = release


def infer_format(data: bytes):
    """iree_hal_cuda_native_executable_infer_format: (format string, inferred size)."""
    fmt = ctypes.create_string_buffer(16)
    size = c_size_t()
    buf = ctypes.create_string_buffer(data, len(data))
    check(lib().iree_b200_executable_infer_format(buf, len(data), 16, fmt, byref(size)))
    return fmt.value.decode(), size.value


class CommandBuffer:
    def __init__(self, device: Device):
        h = c_void_p()
        check(lib().iree_b200_command_buffer_create(device._h, byref(h)))
        self._h = h
        self.device = device

    def begin(self):
        check(lib().iree_b200_command_buffer_begin(self._h))
        return self

    def end(self):
        check(lib().iree_b200_command_buffer_end(self._h))
        return self

    def barrier(self):
        check(lib().iree_b200_command_buffer_execution_barrier(self._h))

    def fill(self, target: BufferRef, pattern: bytes):
        check(lib().iree_b200_command_buffer_fill_buffer(self._h, target, pattern, len(pattern)))

    def update(self, source: bytes, target: BufferRef):
        check(lib().iree_b200_command_buffer_update_buffer(self._h, source, 0, target))

    def copy(self, source: BufferRef, target: BufferRef):
        check(lib().iree_b200_command_buffer_copy_buffer(self._h, source, target))

    def dispatch(self, executable: Executable, ordinal: int, workgroup_count, constants, bindings):
        wg = (c_uint32 * 3)(*workgroup_count)
        cs = (c_uint32 * max(len(constants), 1))(*[c & 0xFFFFFFFF for c in constants])
        bs = (BufferRef * max(len(bindings), 1))(*bindings)
        check(lib().iree_b200_command_buffer_dispatch(self._h, executable._h, ordinal, wg, cs,
                                                     len(constants), bs, len(bindings), 0))

    def collective(self, channel, op: int, param: int, send: BufferRef, recv: BufferRef,
                   element_count: int):
        check(lib().iree_b200_command_buffer_collective(self._h, channel._h, op, param, send, recv,
                                                       element_count))

    def release(self):
        if getattr(self, "_h", None):
            lib().iree_b200_command_buffer_release(self._h)
            self._h = None

    __del__ = release


NULL_REF = BufferRef(None, 0, 0)


class Channel:
    """iree_hal_channel_t over NCCL (include/iree_b200_hal.h, Channels section)."""

    def __init__(self, device: Device = None, rank=-1, count=-1, id_bytes: bytes = None,
                 _handle=None):
        if _handle is not None:
            self._h = _handle
            return
        params = ChannelParams()
        params.flags = 0
        self._id_buf = None
        if id_bytes is not None:
            self._id_buf = ctypes.create_string_buffer(id_bytes, CHANNEL_ID_SIZE)
            params.id = ctypes.cast(self._id_buf, c_void_p)
            params.id_length = CHANNEL_ID_SIZE
        params.rank = rank
        params.count = count
        h = c_void_p()
        check(lib().iree_b200_channel_create(device._h, byref(params), byref(h)))
        self._h = h

    @staticmethod
    def create_id() -> bytes:
        buf = ctypes.create_string_buffer(CHANNEL_ID_SIZE)
        check(lib().iree_b200_channel_id_create(buf))
        return buf.raw

    def attach_symmetric_region(self, executable, byte_length: int, peer_bases, peer_flag_blocks,
                                multicast_base: int = 0, max_blocks: int = 0):
        """In-place all_reduce(sum) inside the region then runs on the NVLink kernels."""
        n = len(peer_bases)
        bases = (c_uint64 * n)(*peer_bases)
        flags = (c_uint64 * n)(*peer_flag_blocks)
        p = SymmetricRegionParams(byte_length, bases, multicast_base, flags, max_blocks)
        check(lib().iree_b200_channel_attach_symmetric_region(self._h, executable._h, byref(p)))

    def collective_status(self) -> int:
        """0, or why a symmetric kernel of this channel gave up (watchdog; see the header)."""
        v = ctypes.c_uint32()
        check(lib().iree_b200_channel_query_collective_status(self._h, byref(v)))
        return v.value

    def collective_counts(self):
        a, b = c_uint64(), c_uint64()
        lib().iree_b200_channel_query_collective_counts(self._h, byref(a), byref(b))
        return {"symmetric": a.value, "nccl": b.value}

    def rank_and_count(self):
        r, c = c_int32(), c_int32()
        lib().iree_b200_channel_query_rank_and_count(self._h, byref(r), byref(c))
        return r.value, c.value

    def split(self, color: int, key: int):
        h = c_void_p()
        check(lib().iree_b200_channel_split(self._h, color, key, 0, byref(h)))
        return Channel(_handle=h) if h.value else None

    def release(self):
        if getattr(self, "_h", None):
            lib().iree_b200_channel_release(self._h)
            self._h = None

    __del__ = release


def default_rank_and_count():
    r, c = c_int32(), c_int32()
    check(lib().iree_b200_channel_provider_query_default_rank_and_count(byref(r), byref(c)))
    return r.value, c.value


def exchange_default_id(rank: int, id_bytes: bytes) -> bytes:
    """iree_hal_channel_provider_exchange_default_id without MPI: rank 0 publishes |id_bytes|
    (file named by IREE_B200_CHANNEL_ID_FILE), every other rank receives them."""
    buf = ctypes.create_string_buffer(id_bytes, CHANNEL_ID_SIZE)
    check(lib().iree_b200_channel_provider_exchange_default_id(rank, buf))
    return buf.raw


class Ops:
    """The dense-contraction dispatches (include/iree_b200_ops.h)."""

    def __init__(self, device: Device, cda1: bytes = None):
        h = c_void_p()
        if cda1 is None:
            check(lib().iree_b200_ops_create(device._h, byref(h)))
        else:
            buf = ctypes.create_string_buffer(cda1, len(cda1))
            check(lib().iree_b200_ops_create_from_data(device._h, buf, len(cda1), byref(h)))
        self._h = h
        self.device = device

    @classmethod
    def offline(cls, sm_count: int = 148, cda1: bytes = None) -> "Ops":
        """A selection-only object (no GPU): select_* and executable lookups work, anything
        that would record or run a kernel raises FAILED_PRECONDITION."""
        self = cls.__new__(cls)
        h = c_void_p()
        if cda1 is None:
            check(lib().iree_b200_ops_create_offline(None, 0, sm_count, byref(h)))
        else:
            buf = ctypes.create_string_buffer(cda1, len(cda1))
            check(lib().iree_b200_ops_create_offline(buf, len(cda1), sm_count, byref(h)))
        self._h = h
        self.device = None
        return self

    @property
    def executable(self) -> Executable:
        return Executable(c_void_p(lib().iree_b200_ops_executable(self._h)), owned=False)

    @property
    def dispatch_count(self) -> int:
        return lib().iree_b200_ops_dispatch_count(self._h)

    def set_reserved_sms(self, reserved_sms: int):
        check(lib().iree_b200_ops_set_reserved_sms(self._h, reserved_sms))

    def set_host_pipeline(self, min_bytes: int, block_m: int = 2048, block_n: int = 2048):
        check(lib().iree_b200_ops_set_host_pipeline(self._h, min_bytes, block_m, block_n))

    def select_matmul(self, element_type, batch, M, N, K, flags=0, alignment=256) -> dict:
        sel = MatmulSelection()
        check(lib().iree_b200_ops_select_matmul(self._h, _etype(element_type), batch, M, N, K,
                                                flags, alignment, byref(sel)))
        return {"export_ordinal": sel.export_ordinal,
                "workgroup_count": list(sel.workgroup_count), "constants": list(sel.constants)}

    def record_matmul(self, cb: CommandBuffer, element_type, batch, M, N, K, lhs: BufferRef,
                      rhs: BufferRef, out: BufferRef, flags=0):
        check(lib().iree_b200_ops_record_matmul(self._h, cb._h, _etype(element_type), batch, M, N,
                                                K, flags, lhs, rhs, out))

    def matmul_host(self, element_type, batch, M, N, K, lhs_ptr: int, rhs_ptr: int, out_ptr: int,
                    flags=0):
        check(lib().iree_b200_ops_matmul_host(self._h, _etype(element_type), batch, M, N, K, flags,
                                              lhs_ptr, rhs_ptr, out_ptr))

    def record_matmul_bias(self, cb: CommandBuffer, element_type, batch, M, N, K, lhs: BufferRef,
                           rhs: BufferRef, bias: BufferRef, out: BufferRef, flags=0):
        """out = act((out +) lhs·rhs + bias[n]) in one dispatch (MATMUL_BIAS implied; MATMUL_RELU,
        MATMUL_TRUNCATE_RESULT optional)."""
        check(lib().iree_b200_ops_record_matmul_bias(self._h, cb._h, _etype(element_type), batch,
                                                     M, N, K, flags, lhs, rhs, bias, out))

    def matmul_bias_host(self, element_type, batch, M, N, K, lhs_ptr: int, rhs_ptr: int,
                         bias_ptr: int, out_ptr: int, flags=0):
        check(lib().iree_b200_ops_matmul_bias_host(self._h, _etype(element_type), batch, M, N, K,
                                                   flags, lhs_ptr, rhs_ptr, bias_ptr, out_ptr))

    def record_matmul_mxfp4(self, cb: CommandBuffer, M, N, K, lhs: BufferRef, lhs_scales: BufferRef,
                            rhs: BufferRef, rhs_scales: BufferRef, out: BufferRef, flags=0):
        """MXFP4 matmul (pending its first GPU run; include/iree_b200_ops.h)."""
        check(lib().iree_b200_ops_record_matmul_mxfp4(self._h, cb._h, M, N, K, flags, lhs,
                                                      lhs_scales, rhs, rhs_scales, out))

    @staticmethod
    def matmul_mxfp4_workspace_size(M, N, K) -> int:
        return lib().iree_b200_ops_matmul_mxfp4_workspace_size(M, N, K)

    @staticmethod
    def matmul_mxfp4_scale_workspace_size(M, N, K) -> int:
        return lib().iree_b200_ops_matmul_mxfp4_scale_workspace_size(M, N, K)

    def record_matmul_mxfp4_tensor(self, cb: CommandBuffer, M, N, K, lhs: BufferRef,
                                   lhs_scales: BufferRef, rhs: BufferRef, rhs_scales: BufferRef,
                                   workspace: BufferRef, out: BufferRef, flags=0):
        """MXFP4 via bf16 expansion + the tcgen05 GEMM (include/iree_b200_ops.h)."""
        check(lib().iree_b200_ops_record_matmul_mxfp4_tensor(
            self._h, cb._h, M, N, K, flags, lhs, lhs_scales, rhs, rhs_scales,
            workspace if workspace is not None else NULL_REF, out))

    def matmul_mxfp4_host(self, M, N, K, lhs_ptr: int, lhs_scales_ptr: int, rhs_ptr: int,
                          rhs_scales_ptr: int, out_ptr: int, flags=0):
        check(lib().iree_b200_ops_matmul_mxfp4_host(self._h, M, N, K, flags, lhs_ptr,
                                                    lhs_scales_ptr, rhs_ptr, rhs_scales_ptr,
                                                    out_ptr))

    def record_mmt4d(self, cb: CommandBuffer, element_type, M1, N1, K1, M0, N0, K0,
                     lhs: BufferRef, rhs: BufferRef, out: BufferRef, flags=0):
        """linalg.mmt4d: lhs [M1,K1,M0,K0], rhs [N1,K1,N0,K0] -> out [M1,N1,M0,N0]."""
        check(lib().iree_b200_ops_record_mmt4d(self._h, cb._h, _etype(element_type), M1, N1, K1,
                                               M0, N0, K0, flags, lhs, rhs, out))

    @staticmethod
    def mmt4d_workspace_size(element_type, M1, N1, K1, M0, N0, K0) -> int:
        return lib().iree_b200_ops_mmt4d_workspace_size(_etype(element_type), M1, N1, K1, M0, N0, K0)

    def record_mmt4d_tensor(self, cb: CommandBuffer, element_type, M1, N1, K1, M0, N0, K0,
                            lhs: BufferRef, rhs: BufferRef, workspace: BufferRef, out: BufferRef,
                            flags=0):
        check(lib().iree_b200_ops_record_mmt4d_tensor(
            self._h, cb._h, _etype(element_type), M1, N1, K1, M0, N0, K0, flags, lhs, rhs,
            workspace if workspace is not None else NULL_REF, out))

    def mmt4d_host(self, element_type, M1, N1, K1, M0, N0, K0, lhs_ptr: int, rhs_ptr: int,
                   out_ptr: int, flags=0):
        check(lib().iree_b200_ops_mmt4d_host(self._h, _etype(element_type), M1, N1, K1, M0, N0, K0,
                                             flags, lhs_ptr, rhs_ptr, out_ptr))

    def select_attention(self, element_type, B, M, K1, K2, N, scale, mask_kind=0, flags=0,
                         alignment=256) -> dict:
        sel = AttentionSelection()
        check(lib().iree_b200_ops_select_attention(self._h, _etype(element_type), B, M, K1, K2, N,
                                                   scale, mask_kind, flags, alignment, byref(sel)))
        return {"workspace_bytes": sel.workspace_bytes,
                "dispatches": [{"export_ordinal": d.export_ordinal,
                                "workgroup_count": list(d.workgroup_count),
                                "constants": list(d.constants)[:d.constant_count]}
                               for d in list(sel.dispatches)[:sel.dispatch_count]]}

    def record_attention(self, cb: CommandBuffer, element_type, B, M, K1, K2, N, scale,
                         q: BufferRef, k: BufferRef, v: BufferRef, out: BufferRef,
                         mask: BufferRef = None, workspace: BufferRef = None, mask_kind=0, flags=0):
        check(lib().iree_b200_ops_record_attention(
            self._h, cb._h, _etype(element_type), B, M, K1, K2, N, scale, mask_kind, flags, q, k, v,
            mask if mask is not None else NULL_REF,
            workspace if workspace is not None else NULL_REF, out))

    def attention_host(self, element_type, B, M, K1, K2, N, scale, q_ptr: int, k_ptr: int,
                       v_ptr: int, out_ptr: int, mask_ptr: int = 0, mask_kind=0, flags=0):
        check(lib().iree_b200_ops_attention_host(self._h, _etype(element_type), B, M, K1, K2, N,
                                                 scale, mask_kind, flags, q_ptr, k_ptr, v_ptr,
                                                 mask_ptr or None, out_ptr))

    def release(self):
        if getattr(self, "_h", None):
            lib().iree_b200_ops_release(self._h)
            self._h = None

    __del__ = release


def _etype(t) -> int:
    return ELEMENT_TYPES[t] if isinstance(t, str) else int(t)
