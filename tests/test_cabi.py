"""CPU suite, part 3: the C-ABI library loads and exports every symbol include/hutoken_b200.h declares; the
product path fails loudly (no CPU fallback) when there is no CUDA device.  No compute calls."""
import ctypes as C
import os
import re
import subprocess
import sys

import pytest

from common import ROOT

HEADER = os.path.join(ROOT, "include", "hutoken_b200.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(hut_[a-z_]+)\s*\(", src)))


def test_library_exports_header_symbols():
    from hutoken_b200 import capi
    lib = capi.load_library()
    names = _declared()
    assert len(names) >= 11
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/hutoken_b200.h but not exported"
    assert sorted(capi.EXPORTS) == names
    out = subprocess.run(["nm", "-D", "--defined-only", capi.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (hut_\w+)", out))
    assert set(names) <= exported


def test_extension_method_table():
    """Same module name and method names as the reference extension (lib.c:1128-1153)."""
    from hutoken_b200 import _hutoken
    for m in ("initialize", "encode", "batch_encode", "decode", "batch_decode", "bpe_train", "bbpe_train"):
        assert callable(getattr(_hutoken, m))
    with pytest.raises(RuntimeError, match="Vocabulary is not initialized for encoding"):
        _hutoken.encode("x")
    with pytest.raises(RuntimeError, match="Vocabulary is not initialized for encoding"):
        _hutoken.batch_encode(["x"], 1)
    with pytest.raises(TypeError):
        _hutoken.initialize("only-one-argument")  # special_file_path is mandatory (lib.c:203)


def test_packed_surface_host_side():
    """pack_texts / batch_encode_packed: host-side packing and argument checks; no device needed."""
    import numpy as np
    import hutoken_b200 as hutoken
    data, offs = hutoken.pack_texts(["ab", "", "őű", b"raw\xff", "cut\0here"])
    assert data == "ab".encode() + "őű".encode() + b"raw\xff" + b"cut"
    assert offs.dtype == np.uint64 and offs.tolist() == [0, 2, 2, 6, 10, 13]
    assert hutoken.pack_texts([])[1].tolist() == [0]
    with pytest.raises(TypeError, match="Expected a list of strings"):
        hutoken.pack_texts("not a list")
    with pytest.raises(TypeError, match="Expected a list of strings"):
        hutoken.pack_texts([1, 2])
    # before initialize(): the reference's message, wrapped like batch_encode's (hutoken.py:134-139)
    with pytest.raises(RuntimeError, match="hutoken: Error encoding texts: Vocabulary is not initialized"):
        hutoken.batch_encode_packed(["x"])
    with pytest.raises(RuntimeError, match="non-decreasing"):
        hutoken.batch_encode_packed(b"abcd", [0, 3, 2])


def test_no_cpu_fallback_without_device():
    """On a box without a GPU hut_create must fail with a CUDA status, not fall back to anything."""
    code = (
        "import sys; sys.path[:0]=[%r, %r]\n"
        "import synth\n"
        "from hutoken_b200 import capi\n"
        "v, s = synth.PATHS['gpt2_small']\n"
        "try:\n"
        "    capi.Engine(v, s, is_byte_encoder=True)\n"
        "except capi.HutError as e:\n"
        "    print('STATUS', e.status)\n"
        "else:\n"
        "    print('CREATED')\n" % (ROOT, os.path.join(ROOT, "tools")))
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env)
    assert "STATUS 4" in out.stdout, out.stdout + out.stderr


def test_product_never_touches_oracle():
    """Nothing under hutoken_b200/ or include/ may reference oracle/ (the checker is test infrastructure)."""
    for base in ("hutoken_b200", "include"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, base)):
            for fn in files:
                if fn.endswith((".py", ".c", ".cu", ".cuh", ".h")) or fn == "Makefile":
                    text = open(os.path.join(dirpath, fn), errors="replace").read()
                    assert "hut_oracle" not in text and "from oracle" not in text and "import oracle" not in text, fn
