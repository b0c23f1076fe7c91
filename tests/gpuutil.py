"""helpers for the -m gpu parity tests: torch is only the device allocator; every compute call goes through the C ABI"""
import ctypes as C

import numpy as np
import torch


# device tensors handed out by dev() stay alive until the end of the test (conftest clears this list): a raw pointer
# taken from a temporary would otherwise dangle as soon as the caching allocator reuses the block
_KEEP = []


def dev(a):
    """numpy -> device tensor (bit-preserving for uint16-viewed fp16)"""
    a = np.ascontiguousarray(a)
    if a.dtype in (np.uint16, np.float16):
        a = a.view(np.int16)
    elif a.dtype == np.uint32:
        a = a.view(np.int32)
    elif a.dtype == np.uint64:
        a = a.view(np.int64)
    t = torch.from_numpy(a).cuda()
    _KEEP.append(t)
    return t


def zeros(n, dtype):
    return torch.zeros(n, dtype=dtype, device="cuda")


def host(t, dtype=None):
    a = t.detach().cpu().numpy()
    return a if dtype is None else a.view(dtype)


def p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def sync():
    torch.cuda.synchronize()
