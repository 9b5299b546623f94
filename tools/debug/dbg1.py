import sys, os
sys.path[:0] = [os.path.join(os.path.dirname(__file__), "..", ".."), os.path.join(os.path.dirname(__file__), ".."), os.path.join(os.path.dirname(__file__), "..", "..", "tests")]
import numpy as np, synth
from common import engine_for, oracle_for, pack, first_diff
name = sys.argv[1] if len(sys.argv) > 1 else "gpt2"
eng, orc = engine_for(name), oracle_for(name)
for texts in (["hello world"], ["hello world", "", "a b c d e f g", "x" * 50, "the the the the"],):
    g, w = eng.batch_encode(texts), orc.batch_encode(texts)
    print(texts[:2], "OK" if g == w else ("DIFF", g, w))
for nb in (2000, 20000, 1 << 20):
    data, offs = synth.corpus(7, nb)
    for rep in range(2):
        ids, tok = eng.encode_packed(data, offs)
        oids, otok = orc.encode_packed(data, offs)
        i = first_diff(ids, oids)
        print(nb, rep, "n", len(ids), len(oids), "first diff", i, ids[max(0,i-3):i+5].tolist(), oids[max(0,i-3):i+5].tolist(),
              "tok ok", np.array_equal(tok, otok), eng.stats())
