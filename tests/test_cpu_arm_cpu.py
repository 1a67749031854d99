"""The CPU-baseline harness (oracle/cpu_arm.py, bench.py --impl reference) on small sizes: it must run the
UNMODIFIED reference classes (kind == "reference" wherever a copy of the reference is present — the live tree here,
the sha256-checked staged copy on the GPU box), produce sane rates, and print exactly one JSON line."""

import json
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT
from oracle import cpu_arm, ref_loader, stage_reference

needs_ref = pytest.mark.skipif(not ref_loader.available(), reason="no copy of the reference on this machine")


def test_staged_copy_is_byte_identical_to_the_reference():
    if not os.path.isdir("/root/reference/agilerl"):
        pytest.skip("reference tree absent (GPU box): the staged copy is checked by its manifest")
    assert stage_reference.stage() == stage_reference.STAGE_ROOT and stage_reference.verify()
    for rel in stage_reference.FILES:
        with open(os.path.join("/root/reference", rel), "rb") as a, open(os.path.join(stage_reference.STAGE_ROOT, rel), "rb") as b:
            assert a.read() == b.read(), rel
    # a tampered copy is refused
    victim = os.path.join(stage_reference.STAGE_ROOT, stage_reference.FILES[0])
    orig = open(victim, "rb").read()
    try:
        open(victim, "ab").write(b"\n# edited\n")
        assert not stage_reference.verify()
    finally:
        open(victim, "wb").write(orig)
    assert stage_reference.verify()


@needs_ref
def test_c2_leg_runs_the_reference_classes():
    ref = ref_loader.load_reference()
    w = cpu_arm.C2Reference(cap=4096, B=32, fill_seconds=1.0)
    assert w.kind == "reference" and type(w.per) is ref.PrioritizedReplayBuffer and type(w.nstep) is ref.MultiStepReplayBuffer
    assert w.per.size == 4096 and w.per.sum_tree.capacity == 4096
    # the direct leaf fill left a consistent tree: every internal node is op(left, right)
    t = np.array(w.per.sum_tree.tree)
    assert np.array_equal(t[1:4096], t[2:8192:2] + t[3:8192:2])
    m = np.array(w.per.min_tree.tree)
    assert np.array_equal(m[1:4096], np.minimum(m[2:8192:2], m[3:8192:2]))
    assert w.agent_step() == 32
    r = cpu_arm.time_c2(0.3, cap=4096, B=32, fill_seconds=0.5)
    assert r["kind"] == "reference" and r["value"] > 0 and r["cores"] == 1 and "unmodified reference" in r["sample"]


@needs_ref
def test_other_legs_run_and_say_what_they_ran():
    for r in (cpu_arm.time_c3(0.2, T=16, N=64, P=2), cpu_arm.time_c4(0.2, cap=1 << 12, B=64),
              cpu_arm.time_uniform(0.2, 5000, 32, cpu_arm.cartpole_row, "small", fill_seconds=0.2),
              cpu_arm.time_insert(0.2, cap=256, num_envs=4)):
        assert r["kind"] == "reference" and r["value"] > 0 and r["sample"]


@needs_ref
def test_bench_reference_arm_prints_one_json_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "0", "--no-also"], capture_output=True, text=True, timeout=600,
                         env={**os.environ, "RANK": "0", "WORLD_SIZE": "1"})
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "per_sampled_transitions_per_s" and d["value"] > 0
    assert d["cpu_baseline"]["kind"] == "reference" and d["e2e"]["value"] == d["value"] and d["gpu_launches"] == 0
    # a non-zero rank under torchrun prints nothing and exits 0
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1"],
                         capture_output=True, text=True, timeout=60, env={**os.environ, "RANK": "1", "WORLD_SIZE": "2"})
    assert out.returncode == 0 and out.stdout.strip() == ""
