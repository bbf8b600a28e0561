"""The allocation-free text entry point against the regular one, through the C ABI.

rfq_text_normalise_noalloc exists for callers that may not allocate GPU memory (a plugin inside a
serving process whose allocator owns the device, a latency-critical handler, a host with a frozen
memory map): its scratch is a fixed array inside the library instead of the engine's growing pool.
It must give byte-identical results to rfq_text_parse — checked here on two calls in a row, since
the second call finds the first call's leftovers in the shared scratch — and refuse oversize input.
This file sorts last on purpose so a `-x` run has already exercised everything else.
"""
import ctypes as C

import pytest

from tests import corpus

pytestmark = pytest.mark.gpu


def _noalloc(lib, torch, doc: bytes, stream_ptr=None) -> bytes:
    n = len(doc)
    cap = max(16, (n + 15) // 16 * 16)
    d_in = torch.zeros(cap, dtype=torch.uint8, device="cuda:0")
    if n:
        d_in[:n] = torch.frombuffer(bytearray(doc), dtype=torch.uint8).to("cuda:0")
    d_out = torch.full((cap,), 0xAA, dtype=torch.uint8, device="cuda:0")
    d_len = torch.full((1,), -1, dtype=torch.int64, device="cuda:0")
    torch.cuda.synchronize()
    rc = lib.rfq_text_normalise_noalloc(0, d_in.data_ptr(), n, d_out.data_ptr(), d_len.data_ptr(),
                                        stream_ptr)
    assert rc == 0, rc
    torch.cuda.synchronize()
    out_len = int(d_len.item())
    assert 0 <= out_len <= n
    return bytes(d_out[:out_len].cpu().numpy().tobytes())


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    import torch
    from replisense_rfq_b200 import Engine, _lib, build
    build.build()
    lib = _lib.load()
    eng = Engine(0)
    # two different documents back to back (the second sees the first one's scratch contents),
    # then the first again, then edge sizes
    big = ("  \r\n" + "line ₹ 1,200 qty: 750\r\n" * 40000 + "\t \n").encode() + b"\xff\xfe"
    docs = [corpus.text_body(11), big, corpus.text_body(11), b"", b" \n ", b"x",
            b"a" * 512, ("é" * 1000).encode(), b"\xe2\x82" * 300 + b"z"]
    docs += [corpus.text_body(s) for s in range(300, 330)]
    want = [eng.text_batch([d]).doc_bytes(0) for d in docs]   # existing entry point, C ABI
    got = [_noalloc(lib, torch, d) for d in docs]
    assert got == want
    # and against the oracle directly, so this test stands on its own
    from oracle import file_parser_oracle as O
    assert [g.decode("utf-8") for g in got] == [O.parse_txt_bytes(d) for d in docs]
    # a second full round: identical again
    assert [_noalloc(lib, torch, d) for d in docs[:4]] == want[:4]
    # on an explicit stream too
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        assert _noalloc(lib, torch, docs[1], st.cuda_stream) == want[1]


def test_noalloc_rejects_oversize_and_bad_arguments():
    import torch
    from replisense_rfq_b200 import _lib, build
    build.build()
    lib = _lib.load()
    buf = torch.zeros(64, dtype=torch.uint8, device="cuda:0")
    n64 = torch.zeros(1, dtype=torch.int64, device="cuda:0")
    too_big = _lib.NOALLOC_MAX_BYTES + 1
    assert lib.rfq_text_normalise_noalloc(0, buf.data_ptr(), too_big, buf.data_ptr(), n64.data_ptr(),
                                          None) == _lib.ERR_TOO_LARGE
    assert lib.rfq_text_normalise_noalloc(0, None, 8, buf.data_ptr(), n64.data_ptr(), None) == -2
    assert lib.rfq_text_normalise_noalloc(0, buf.data_ptr() + 1, 8, buf.data_ptr(), n64.data_ptr(),
                                          None) == -2
    # exactly at the limit is accepted
    n = _lib.NOALLOC_MAX_BYTES
    d_in = torch.full((n,), ord("a"), dtype=torch.uint8, device="cuda:0")
    d_out = torch.zeros(n, dtype=torch.uint8, device="cuda:0")
    assert lib.rfq_text_normalise_noalloc(0, d_in.data_ptr(), n, d_out.data_ptr(), n64.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert int(n64.item()) == n and bool((d_out == ord("a")).all())
