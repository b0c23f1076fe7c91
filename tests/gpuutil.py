"""helpers for the -m gpu parity tests: torch is only the device allocator; every compute call goes through the C ABI"""
import ctypes as C

import numpy as np
import torch


def dev(a):
    """numpy -> device tensor (bit-preserving for uint16-viewed fp16)"""
    a = np.ascontiguousarray(a)
    if a.dtype == np.uint16:
        return torch.from_numpy(a.view(np.int16)).cuda()
    if a.dtype == np.uint32:
        return torch.from_numpy(a.view(np.int32)).cuda()
    if a.dtype == np.uint64:
        return torch.from_numpy(a.view(np.int64)).cuda()
    return torch.from_numpy(a).cuda()


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
