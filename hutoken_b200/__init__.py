"""hutoken_b200 -- B200-native BPE encode engine behind hutoken's initialize / encode / batch_encode surface.

    import hutoken_b200 as hutoken
    hutoken.initialize("vocab.txt", "special_chars.txt", is_byte_encoder=True)
    ids = hutoken.encode("hello world")
    batches = hutoken.batch_encode(["a", "b"], num_threads=8)
    ids, tok_offsets = hutoken.batch_encode_packed(["a", "b"])   # int32 / uint64 ndarrays, no PyLong per id

Every id is produced by sm_100a CUDA kernels in libhutoken_b200.so (C-ABI: include/hutoken_b200.h).
There is no CPU fallback: importing works anywhere, but initialize() fails without a CUDA device.
"""
from .hutoken import (batch_decode, batch_encode, batch_encode_packed, bbpe_train, bpe_train,  # noqa: F401
                      decode, encode, initialize, initialize_foma, look_up_word, pack_texts)

__all__ = ["initialize", "encode", "batch_encode", "batch_encode_packed", "pack_texts", "decode", "batch_decode", "bpe_train", "bbpe_train",
           "initialize_foma", "look_up_word"]
