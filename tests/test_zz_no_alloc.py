"""hut_decode_device_noalloc (include/hutoken_b200.h): the token-wise decode kernels with their scratch in a file-scope
device array of the library and caller-owned outputs -- no cudaMalloc on the call.  No counterpart in the reference.
Checked against the existing entry points (hut_decode_batch /
hut_decode_device) byte for byte, on calls in a row that share the array.

Collected last on purpose (the file name).
"""
import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu


def _batches(eng, seed):
    """ids of a corpus encoded on the GPU (unknown symbols dropped) cut into documents three different ways"""
    data, offs = synth.corpus(seed, 3 << 20, special_every=180)
    ids, _ = eng.encode_packed(data, offs)
    ids = np.ascontiguousarray(ids[ids >= 0])
    rng = np.random.default_rng(seed)
    n = len(ids)
    out = []
    for n_use, n_cuts in ((n, 700), (n // 5, 9000), (n, 0)):  # ragged; many empty / tiny documents; one document
        cuts = np.sort(rng.integers(0, n_use + 1, size=n_cuts))
        tok = np.concatenate([[0], cuts, [n_use]]).astype(np.uint64)
        out.append((ids[:n_use], tok))
    return out


@pytest.mark.parametrize("name", ["gpt2_small", "llama_small", "gpt2"])
def test_no_alloc_decode_equals_decode_device(name):
    import torch
    from hutoken_b200 import capi
    dev = torch.device("cuda:0")
    v, s = synth.PATHS[name]
    eng = capi.Engine(v, s, **synth.MODEL_KW[name])
    # before the decode tables exist the call says so (it may not build them: that allocates)
    z = torch.zeros(8, dtype=torch.int64, device=dev)
    with pytest.raises(capi.HutError, match="decode tables not built"):
        eng.decode_device_noalloc(z.data_ptr(), 1, z.data_ptr(), 1, z.data_ptr(), 8, z.data_ptr())
    batches = _batches(eng, 91)
    want = [eng.decode_packed(ids, tok) for ids, tok in batches]  # existing entry point; also builds the tables
    assert eng.decode_stats()["path"] == "token-wise"
    cap = max(len(t) for t, _ in want) + 64
    d_text = torch.empty(cap, dtype=torch.uint8, device=dev)
    d_toff = torch.empty(max(len(tok) for _, tok in batches) + 8, dtype=torch.int64, device=dev)
    dev_in = [(torch.from_numpy(ids).to(dev), torch.from_numpy(tok.astype(np.int64)).to(dev)) for ids, tok in batches]
    torch.cuda.synchronize()
    free0 = None
    # calls in a row on the shared array: large, small (stale lengths / first-document entries of the large one are
    # still in the array), one document, then the first again
    for rnd, b in enumerate((0, 1, 2, 0, 1)):
        (ids, tok), (w_text, w_off), (d_ids, d_tok) = batches[b], want[b], dev_in[b]
        d_text.fill_(0xAA)
        d_toff.fill_(-1)
        torch.cuda.synchronize()
        if rnd == 3:
            free0 = torch.cuda.mem_get_info(dev)[0]
        n = eng.decode_device_noalloc(d_ids.data_ptr(), len(ids), d_tok.data_ptr(), len(tok) - 1, d_text.data_ptr(), cap,
                                      d_toff.data_ptr())
        if rnd == 3:  # kernels loaded by now: the call itself takes no device memory
            assert torch.cuda.mem_get_info(dev)[0] == free0
        assert n == len(w_text), (name, rnd)
        got_off = d_toff[:len(tok)].cpu().numpy().astype(np.uint64)
        assert np.array_equal(got_off, w_off), (name, rnd)
        got = d_text.cpu().numpy()
        assert got[:n].tobytes() == w_text.tobytes(), (name, rnd)
        assert (got[n:] == 0xAA).all(), (name, rnd, "wrote past the text")
        assert int(d_toff[len(tok)].item()) == -1
    # the same ids / offsets through hut_decode_device (device-resident outputs owned by the context)
    (ids, tok), (d_ids, d_tok) = batches[0], dev_in[0]
    t_ptr, o_ptr, n2 = eng.decode_device(d_ids.data_ptr(), len(ids), d_tok.data_ptr(), len(tok) - 1)
    n = eng.decode_device_noalloc(d_ids.data_ptr(), len(ids), d_tok.data_ptr(), len(tok) - 1, d_text.data_ptr(), cap,
                                  d_toff.data_ptr())
    assert n == n2

    class View:
        def __init__(self, ptr, nbytes):
            self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}

    assert torch.equal(torch.as_tensor(View(t_ptr, n2), device=dev), d_text[:n])
    assert torch.equal(torch.as_tensor(View(o_ptr, 8 * len(tok)), device=dev), d_toff[:len(tok)].view(torch.uint8))
    # ---- limits and refusals ----
    with pytest.raises(ValueError, match="at most 16777216 ids"):
        eng.decode_device_noalloc(d_ids.data_ptr(), capi.NOALLOC_MAX_IDS + 1, d_tok.data_ptr(), 1, d_text.data_ptr(), cap,
                                  d_toff.data_ptr())
    with pytest.raises(capi.HutError, match="text_cap too small"):
        eng.decode_device_noalloc(d_ids.data_ptr(), len(ids), d_tok.data_ptr(), len(tok) - 1, d_text.data_ptr(), 100,
                                  d_toff.data_ptr())
    bad = d_ids.clone()
    bad[len(ids) // 2] = eng.vocab_lines
    with pytest.raises(ValueError, match="less than vocab size"):
        eng.decode_device_noalloc(bad.data_ptr(), len(ids), d_tok.data_ptr(), len(tok) - 1, d_text.data_ptr(), cap,
                                  d_toff.data_ptr())
    # empty batch / empty documents
    zt = torch.zeros(4, dtype=torch.int64, device=dev)
    d_toff.fill_(-1)
    assert eng.decode_device_noalloc(d_ids.data_ptr(), 0, zt.data_ptr(), 3, d_text.data_ptr(), cap, d_toff.data_ptr()) == 0
    assert d_toff[:5].tolist() == [0, 0, 0, 0, -1]
    # and a good call still works after the refused ones
    n = eng.decode_device_noalloc(d_ids.data_ptr(), len(ids), d_tok.data_ptr(), len(tok) - 1, d_text.data_ptr(), cap,
                                  d_toff.data_ptr())
    assert d_text[:n].cpu().numpy().tobytes() == want[0][0].tobytes()
    eng.close()
