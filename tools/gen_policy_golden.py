#!/usr/bin/env python3
"""Generate tests/golden/extracted_policy.npz by importing the REFERENCE'S OWN src/extracted_policy.py, unmodified.

Runs only in the build container (needs /root/reference).  The fixture holds
  * the reference's known-answer test (extracted_policy.py:480-485): `kat_obs` [1, 66] -> `kat_expected` [1, 28];
  * the trained weights the reference embeds (W0 [66,256], B0, W2 [256,128], B2, WA [128,28], BA), stored as float32
    (they are float32 values printed with 8-9 significant digits);
  * outputs of ExtractedPolicy.act (float64 numpy: matmul + tanh, :471-478) on 512 seeded observations (`obs`,
    `act`), with the float32-rounded weights, so the CUDA path and the oracle see bit-identical parameters.

Usage: python tools/gen_policy_golden.py
"""
import os
import sys

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("DM_REFERENCE", "/root/reference")
sys.path.insert(0, os.path.join(REF, "src"))

import extracted_policy as ref  # noqa: E402  (reference)

pol = ref.ExtractedPolicy()
pol.test()  # the reference's own assertion; its observation literal is recovered below by intercepting act()
captured = {}
orig_act = pol.act


def spy(obs):
    captured["obs"] = np.array(obs, dtype=np.float64)
    out = orig_act(obs)
    captured["out"] = np.array(out, dtype=np.float64)
    return out


pol.act = spy
pol.test()
pol.act = orig_act
kat_obs, kat_out = captured["obs"], captured["out"]

W = {k: np.asarray(getattr(pol, k), dtype=np.float64) for k in ("W0", "B0", "W2", "B2", "WA", "BA")}
W32 = {k: v.astype(np.float32) for k, v in W.items()}
assert all(np.abs(W32[k].astype(np.float64) - W[k]).max() <= 6e-8 * max(1.0, np.abs(W[k]).max()) for k in W)
for k, v in W32.items():  # run the reference on the float32-rounded parameters
    setattr(pol, k, v.astype(np.float64))
rng = np.random.default_rng(0)
obs = np.concatenate([rng.uniform(0, 1, (256, 66)), rng.normal(0, 2.0, (256, 66))]).astype(np.float32)
act = pol.act(obs.astype(np.float64))
kat32 = pol.act(kat_obs.astype(np.float32).astype(np.float64))
# the reference's stated expectation (np.allclose) must survive the float32 rounding of weights and input
assert np.allclose(kat32, kat_out, rtol=1e-5, atol=1e-6), np.abs(kat32 - kat_out).max()
out = os.path.join(REPO, "tests", "golden", "extracted_policy.npz")
np.savez_compressed(out, kat_obs=kat_obs, kat_expected=kat_out, obs=obs, act=act, **W32)
print("wrote", out, os.path.getsize(out), "bytes; KAT drift from float32 rounding: %.2e" % np.abs(kat32 - kat_out).max())
