"""fdls_sort_pairs_static: the allocation-free variant of fdls_sort_pairs (scratch in a static device array inside the
library; include/fdls_abi.h).  Collected last on purpose (file name).

Checked against the existing entry point (caller-provided scratch) through the C-ABI, bit-exact (the sort is stable, so
keys AND values must agree), on two calls in a row that share the static array: the second call has a different size,
different key bits and a device-side element count, so stale contents of the first call would show."""
import ctypes as C
import importlib
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

MAX_PAIRS = 2097152  # FDLS_SORT_STATIC_MAX_PAIRS


def _lib():
    return importlib.import_module("4dlangsplat_b200._lib").load()


def _keys(n, end_bit, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    lo = torch.randint(0, 2 ** 31 - 1, (n,), generator=g, dtype=torch.int64)
    hi = torch.randint(0, 2 ** 31 - 1, (n,), generator=g, dtype=torch.int64)
    keys = (hi << 32) | lo
    if end_bit < 64:
        keys = keys & ((1 << end_bit) - 1)
    keys[: n // 3] = keys[n // 2: n // 2 + n // 3].clone()  # duplicates: stability matters
    return keys.cuda()


def _run(lib, keys, n_max, n_dev, end_bit, static):
    vals = torch.arange(n_max, dtype=torch.int32, device="cuda")
    ka, kb, va, vb = keys.clone(), torch.zeros_like(keys), vals.clone(), torch.zeros_like(vals)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    nd = n_dev.data_ptr() if n_dev is not None else None
    if static:
        sel = lib.fdls_sort_pairs_static(ka.data_ptr(), kb.data_ptr(), va.data_ptr(), vb.data_ptr(), n_max, nd, 0, end_bit,
                                         stream)
    else:
        sb = lib.fdls_sort_scratch_bytes(n_max)
        scratch = torch.empty(sb, dtype=torch.uint8, device="cuda")
        sel = lib.fdls_sort_pairs(ka.data_ptr(), kb.data_ptr(), va.data_ptr(), vb.data_ptr(), n_max, nd, 0, end_bit,
                                  scratch.data_ptr(), sb, stream)
    torch.cuda.synchronize()
    assert sel in (0, 1), "sort failed: %d" % sel
    return (kb, vb) if sel else (ka, va)


@pytest.mark.gpu
def test_sort_pairs_static_matches_sort_pairs_twice_in_a_row():
    lib = _lib()
    # call 1: above 2^20 elements (8 items per thread), all 64 key bits, host-side count
    n1 = 1_500_003
    k1 = _keys(n1, 64, 1)
    want_k, want_v = _run(lib, k1, n1, None, 64, static=False)
    got_k, got_v = _run(lib, k1, n1, None, 64, static=True)
    assert torch.equal(got_k, want_k) and torch.equal(got_v, want_v)
    ref_k, order = torch.sort(k1, stable=True)  # keys < 2^63: signed order == unsigned order
    assert torch.equal(got_k, ref_k) and torch.equal(got_v.long(), order)
    # call 2, straight after, same static array: smaller (4 items per thread), 37 key bits = 5 passes (odd: result in
    # the other buffer), element count read on the device
    n_max2, n2 = 300_000, 123_457
    k2 = _keys(n_max2, 37, 2)
    n_dev = torch.tensor([n2], dtype=torch.int32, device="cuda")
    want_k, want_v = _run(lib, k2, n_max2, n_dev, 37, static=False)
    got_k, got_v = _run(lib, k2, n_max2, n_dev, 37, static=True)
    assert torch.equal(got_k[:n2], want_k[:n2]) and torch.equal(got_v[:n2], want_v[:n2])
    ref_k, order = torch.sort(k2[:n2], stable=True)
    assert torch.equal(got_k[:n2], ref_k) and torch.equal(got_v[:n2].long(), order)
    # and call 1's input again: the array now holds call 2's state
    got_k, got_v = _run(lib, k1, n1, None, 64, static=True)
    ref_k, order = torch.sort(k1, stable=True)
    assert torch.equal(got_k, ref_k) and torch.equal(got_v.long(), order)


@pytest.mark.gpu
def test_sort_pairs_static_limits():
    lib = _lib()
    n = MAX_PAIRS  # exactly the maximum is accepted
    k = _keys(n, 40, 3)
    got_k, got_v = _run(lib, k, n, None, 40, static=True)
    ref_k, order = torch.sort(k, stable=True)
    assert torch.equal(got_k, ref_k) and torch.equal(got_v.long(), order)
    one = torch.zeros(8, dtype=torch.int64, device="cuda")
    rc = lib.fdls_sort_pairs_static(one.data_ptr(), one.data_ptr(), one.data_ptr(), one.data_ptr(), MAX_PAIRS + 1, None, 0,
                                    32, None)
    assert rc == -5 and b"maximum" in lib.fdls_strerror(rc)
    assert lib.fdls_sort_pairs_static(None, None, None, None, 10, None, 0, 32, None) == -1
    assert lib.fdls_sort_pairs_static(None, None, None, None, 0, None, 0, 32, None) in (0, 1)  # empty input: no work


def test_static_limit_matches_header():
    hdr = open(os.path.join(ROOT, "include", "fdls_abi.h")).read()
    assert "#define FDLS_SORT_STATIC_MAX_PAIRS %d" % MAX_PAIRS in hdr
    lib = _lib()
    assert lib.fdls_sort_pairs_static(1, 1, 1, 1, MAX_PAIRS + 1, None, 0, 32, None) == -5  # rejected before any CUDA call
