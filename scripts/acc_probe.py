"""How does tcgen05.mma (kind::f16, bf16 operands) round its fp32 accumulation?
    python scripts/acc_probe.py  -> gpurun_out/umma_acc_probe.json
Every probe row is a 64-long dot product a . b with b = 1: values inside one group of 16 k are added
within ONE instruction, different groups by the accumulate of consecutive instructions.  u = ulp(1.5) = 2^-23.
The observed sum is compared with the round-to-nearest (RN), toward-zero (RZ), down (RD) and up (RU)
roundings of the exact sum."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from vcm_ts_b200 import _probe  # noqa: E402

L = _probe.lib()
u = 2.0 ** -23
rows = []   # (name, {k: value})


def row(name, **kv):
    rows.append((name, {int(k[1:]): v for k, v in kv.items()}))


# across instructions (k = 0 in instruction 0, k = 16 in instruction 1, ...)
row("acc: 1.5 + 0.75u", k0=1.5, k16=0.75 * u)
row("acc: 1.5 + 0.25u", k0=1.5, k16=0.25 * u)
row("acc: 1.5 - 0.25u", k0=1.5, k16=-0.25 * u)
row("acc: 1.5 - 0.75u", k0=1.5, k16=-0.75 * u)
row("acc: -1.5 + 0.25u", k0=-1.5, k16=0.25 * u)
row("acc: -1.5 - 0.25u", k0=-1.5, k16=-0.25 * u)
row("acc: -1.5 - 0.75u", k0=-1.5, k16=-0.75 * u)
row("acc: 1.5 + 0.5u (tie)", k0=1.5, k16=0.5 * u)
row("acc: 1.5+u + 0.5u (tie, odd)", k0=1.5, k1=u, k16=0.5 * u)
row("acc: 1.5 + 0.75u + 0.75u + 0.75u", k0=1.5, k16=0.75 * u, k32=0.75 * u, k48=0.75 * u)
# inside one instruction
row("in: 1.5 + 0.75u", k0=1.5, k1=0.75 * u)
row("in: 1.5 + 0.25u", k0=1.5, k1=0.25 * u)
row("in: 1.5 - 0.25u", k0=1.5, k1=-0.25 * u)
row("in: -1.5 + 0.25u", k0=-1.5, k1=0.25 * u)
row("in: 1.5 + 15 x u/16", **{"k0": 1.5, **{f"k{j}": u / 16 for j in range(1, 16)}})
row("in: 1.5 + 15 x u/64", **{"k0": 1.5, **{f"k{j}": u / 64 for j in range(1, 16)}})
row("in: 1.5 + 8 x u/8 - 0", **{"k0": 1.5, **{f"k{j}": u / 8 for j in range(1, 9)}})
row("in: 1024 + 15 x 2^-14", **{"k0": 1024.0, **{f"k{j}": 2.0 ** -14 for j in range(1, 16)}})
# second instruction's products summed first, then added to the accumulator?
row("acc: 1.5 + (8 x u/8 in one instruction)", **{"k0": 1.5, **{f"k{j}": u / 8 for j in range(16, 24)}})
row("acc: 1.5 + (0.5u + 0.25u in one instruction)", k0=1.5, k16=0.5 * u, k17=0.25 * u)

a = np.zeros((128, 64), dtype=np.float64)
for i, (_, kv) in enumerate(rows):
    for k, v in kv.items():
        a[i, k] = v
at = torch.tensor(a, dtype=torch.float32, device="cuda")
assert torch.equal(at.bfloat16().float(), at), "probe values must be bf16-exact"
b = torch.ones(64, device="cuda")
out = torch.zeros(128, device="cuda")
assert L.vcm_umma_acc_probe(at.data_ptr(), b.data_ptr(), out.data_ptr(), None) == 0
torch.cuda.synchronize()
got = out.cpu().numpy().astype(np.float64)


def roundings(x):
    f = np.float32(x)            # RN
    lo = f if float(f) <= x else np.nextafter(f, np.float32(-np.inf))
    hi = f if float(f) >= x else np.nextafter(f, np.float32(np.inf))
    rz = lo if x >= 0 else hi
    return {"RN": float(f), "RZ": float(rz), "RD": float(lo), "RU": float(hi)}


res = []
score = {"RN": 0, "RZ": 0, "RD": 0, "RU": 0}
for i, (name, kv) in enumerate(rows):
    exact = float(sum(kv.values()))
    r = roundings(exact)
    match = [m for m, v in r.items() if v == got[i]]
    for m in match:
        score[m] += 1
    res.append({"probe": name, "exact_minus_1.5_in_ulp": (abs(exact) - 1.5) / u if abs(exact) < 8 else None,
                "got_minus_1.5_in_ulp": (abs(got[i]) - 1.5) / u if abs(got[i]) < 8 else got[i], "matches": match})
    print(f"{name:48s} exact {exact!r:26} got {got[i]!r:22} matches {match}")
print("rows consistent with each rounding:", score, "of", len(rows))
os.makedirs("gpurun_out", exist_ok=True)
json.dump({"rows": res, "score": score, "n": len(rows)}, open("gpurun_out/umma_acc_probe.json", "w"), indent=1)
