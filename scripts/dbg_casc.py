import sys, numpy as np
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import gpu_common as g, datagen, oracle_lib as o
c = datagen.sorted_ints(1024, 12345).astype(np.int32).view(np.uint8)
co = g.L.CascadedOpts(4096, 4, 1, 1, 1)
comp, dec, status, sizes, actual = g.roundtrip("cascaded", [c], opts=co)
exp = o.casc_compress(c, 4, 1, 1, 1)
a = np.frombuffer(comp[0].tobytes(), np.uint32); b = np.frombuffer(exp.tobytes(), np.uint32)
print("ours", len(a), [hex(x) for x in a[:12]])
print("exp ", len(b), [hex(x) for x in b[:12]])
n = min(len(a), len(b)); d = np.nonzero(a[:n] != b[:n])[0]
print("first diffs", d[:20])
# second array header
oc = a[3]; ec = b[3]
print("ours vals hdr", [hex(x) for x in a[6 + (oc + 3) // 4: 6 + (oc + 3) // 4 + 3]], "exp", [hex(x) for x in b[6 + (ec + 3) // 4: 6 + (ec + 3) // 4 + 3]])
