"""The allocation-free log-prob entry point, driven through the C ABI exactly as an embedding server would.

Some users of this library run inside a process that has already given almost all GPU memory to its own
pool (an inference / scoring server with a pre-sized KV cache) or under a memory budget fixed ahead of
time, so they cannot let a library call cudaMalloc for a work area.  ``dla_lmhead_logprob_noalloc`` keeps
its tile partials in a fixed array inside the library image instead.  This test calls it with raw pointers
via ctypes — no Python wrapper of ours in between — and checks it against the regular entry point
(caller-provided scratch) twice in a row with different shapes, so stale contents left in the shared
array by the first call would show up in the second.  It also checks that oversized requests are refused.

The file name sorts last on purpose: a run that stops at the first failure has then already executed
every other test.
"""
import ctypes
from ctypes import c_int, c_int64, c_void_p

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from dla_b200.ops import build, native

    if not build.is_current():
        build.build()
    lib = ctypes.CDLL(str(native.lib_path()))
    lib.dla_last_error.restype = ctypes.c_char_p
    lib.dla_lmhead_scratch_bytes.restype = ctypes.c_size_t
    lib.dla_lmhead_scratch_bytes.argtypes = [c_int, c_int]
    sig_tail = [c_void_p, c_void_p, c_void_p]
    lib.dla_lmhead_logprob.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64, c_int64, c_void_p] + sig_tail
    lib.dla_lmhead_logprob_noalloc.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64, c_int64] + sig_tail
    for f in (lib.dla_lmhead_logprob, lib.dla_lmhead_logprob_noalloc, lib.dla_set_device):
        f.restype = c_int
    assert lib.dla_set_device(torch.cuda.current_device()) == 0, lib.dla_last_error()
    return lib


def _case(lib, M, V, K, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(M, K, device="cuda", generator=g).to(torch.bfloat16)
    w = (torch.randn(V, K, device="cuda", generator=g) * 0.05).to(torch.bfloat16)
    labels = torch.randint(0, V, (M,), device="cuda", generator=g)
    labels[::5] = -100                                   # ignored rows never write their "picked" slot
    stream = torch.cuda.current_stream().cuda_stream
    # existing entry point: the caller brings the scratch
    scratch = torch.empty(int(lib.dla_lmhead_scratch_bytes(M, V)), dtype=torch.uint8, device="cuda")
    lp_a, lse_a = torch.empty(M, device="cuda"), torch.empty(M, device="cuda")
    rc = lib.dla_lmhead_logprob(x.data_ptr(), w.data_ptr(), labels.data_ptr(), M, V, K, K, K, scratch.data_ptr(),
                                lp_a.data_ptr(), lse_a.data_ptr(), stream)
    assert rc == 0, lib.dla_last_error()
    # new entry point: inputs and outputs only
    lp_b, lse_b = torch.full((M,), 7.0, device="cuda"), torch.full((M,), 7.0, device="cuda")
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]           # the driver's view: sees a cudaMalloc made by ANY library
    rc = lib.dla_lmhead_logprob_noalloc(x.data_ptr(), w.data_ptr(), labels.data_ptr(), M, V, K, K, K,
                                        lp_b.data_ptr(), lse_b.data_ptr(), stream)
    assert rc == 0, lib.dla_last_error()
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before, "the allocation-free entry point changed the device's free memory"
    # same kernel, same inputs: agreement well inside the 3e-2 the log-prob tests allow against fp32
    assert torch.allclose(lp_b, lp_a, atol=1e-4, rtol=1e-5), (lp_b - lp_a).abs().max()
    assert torch.allclose(lse_b, lse_a, atol=1e-4, rtol=1e-5), (lse_b - lse_a).abs().max()
    # and both agree with plain PyTorch fp32 within the suite's tolerance
    logits = x.float() @ w.float().t()
    ref_lse = torch.logsumexp(logits, -1)
    ref_lp = torch.where(labels >= 0, logits.gather(-1, labels.clamp(min=0)[:, None])[:, 0] - ref_lse, torch.zeros_like(ref_lse))
    assert torch.allclose(lp_b, ref_lp, atol=3e-2, rtol=2e-2)
    assert torch.allclose(lse_b, ref_lse, atol=3e-2, rtol=2e-2)


def test_noalloc_logprob_matches_regular_entry_point_twice_in_a_row():
    lib = _lib()
    _case(lib, M=1000, V=5000, K=256, seed=1)     # fills most rows and 20 vocabulary tiles of the shared area
    _case(lib, M=300, V=1000, K=512, seed=2)      # smaller second call on top of the first call's leftovers


def test_noalloc_logprob_rejects_sizes_beyond_the_built_in_area():
    lib = _lib()
    lib.dla_lmhead_logprob_noalloc_max_rows.restype = c_int
    lib.dla_lmhead_logprob_noalloc_max_vocab.restype = c_int
    max_rows, max_vocab = lib.dla_lmhead_logprob_noalloc_max_rows(), lib.dla_lmhead_logprob_noalloc_max_vocab()
    assert max_rows == 1024 and max_vocab == 131072
    x = torch.zeros(8, 64, device="cuda", dtype=torch.bfloat16)
    lab = torch.zeros(max_rows + 8, device="cuda", dtype=torch.int64)
    out = torch.zeros(max_rows + 8, device="cuda")
    rc = lib.dla_lmhead_logprob_noalloc(x.data_ptr(), x.data_ptr(), lab.data_ptr(), max_rows + 1, 512, 64, 64, 64,
                                        out.data_ptr(), out.data_ptr(), None)
    assert rc == -4 and b"exceed" in lib.dla_last_error()
    rc = lib.dla_lmhead_logprob_noalloc(x.data_ptr(), x.data_ptr(), lab.data_ptr(), 8, max_vocab + 1, 64, 64, 64,
                                        out.data_ptr(), out.data_ptr(), None)
    assert rc == -4 and b"vocabulary tiles" in lib.dla_last_error()
