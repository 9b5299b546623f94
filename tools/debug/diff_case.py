"""debug helper: encode a few adversarial strings on the GPU and print where the ids diverge from the oracle"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "tests")]
import numpy as np
from common import engine_for, oracle_for, pack

name = sys.argv[1] if len(sys.argv) > 1 else "gpt2_small"
eng, orc = engine_for(name), oracle_for(name)
cases = {
    "w17": [("w" * 17 + " ") * 500],
    "w17x": ["x" * 100 + ("w" * 17 + " ") * 120],
    "w10": [("w" * 10 + " ") * 500],
    "w20": [("w" * 20 + " ") * 300],
}
for key, docs in cases.items():
    data, offs = pack(docs)
    ids, tok = eng.encode_packed(data, offs)
    oids, otok = orc.encode_packed(data, offs)
    if len(ids) == len(oids) and np.array_equal(ids, oids):
        print(key, "ok", eng.stats())
        continue
    n = min(len(ids), len(oids))
    bad = np.nonzero(ids[:n] != oids[:n])[0]
    # token index -> byte offset via per-pretoken oracle encode
    text = docs[0]
    pre = orc.pretokens(text)
    tpos, bpos, t = [], 0, 0
    for p in pre:
        k = len(orc.encode(p)) if bpos else len(orc.encode(p))
        tpos.append((t, bpos, len(p)))
        t += k
        bpos += len(p)
    def where(ti):
        for (t0, b0, l) in reversed(tpos):
            if t0 <= ti:
                return (t0, b0, l)
    print(key, "DIFF", len(ids), len(oids), "first bad", bad[:10].tolist(), "n bad", len(bad), eng.stats())
    for ti in bad[:6]:
        t0, b0, l = where(int(ti))
        print("   tok", int(ti), "got", int(ids[ti]), "want", int(oids[ti]), "pretoken tok0", t0, "byte", b0, "len", l,
              "byte%960", b0 % 960, "chunk", (b0 % 960) // 32)
