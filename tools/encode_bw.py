"""Dev tool: bandwidth of the resident-record -> blocked bf16 encode at batch 4096."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from deepgo_b200.host import engine
B = 4096
ctx = engine.Context(0)
m = engine.Model(ctx, 3, 64, engine.DG_PREC_BF16, max_batch=B)
b = bench.synthetic_batch(5, B)
m.set_batch_raw(b["raw"], b["player"], b["rank"])
for _ in range(5):
    m.encode_resident()
ctx.prof_begin()
for _ in range(20):
    m.encode_resident()
pe = ctx.prof_end()["encode"]
ms = pe[0] / pe[1]
alg = B * (9 * 361 + 3 + 6 * 361 * 16)
print("encode batch %d: %.1f us, %.0f GB/s algorithmic (%.2f of %.0f GB/s HBM peak)" % (B, ms * 1e3, alg / ms / 1e6, alg / ms / 1e6 / bench.peaks()[2], bench.peaks()[2]))
