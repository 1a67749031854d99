"""TEST INFRASTRUCTURE — generates tests/golden/*.npz by running the UNMODIFIED reference.

Run in the build container (needs /root/reference):  ``python oracle/make_golden.py``
The fixtures are committed; the GPU box (no reference tree) only reads them.
Every fixture stores the inputs and the reference's outputs for one seeded scenario.
"""

from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from oracle import ref_loader  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")


def save(name, **arrays):
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **arrays)
    print("wrote", name, {k: np.asarray(v).shape for k, v in arrays.items()})


def gen_gae():
    for seed, (T, N) in [(0, (16, 8)), (1, (128, 64)), (2, (37, 33))]:
        g = torch.Generator().manual_seed(seed)
        r = torch.randn(T, N, generator=g).numpy()
        v = torch.randn(T, N, generator=g).numpy()
        d = (torch.rand(T, N, generator=g) < 0.05).numpy()
        lv = torch.randn(N, generator=g).numpy()
        ld = (torch.rand(N, generator=g) < 0.1).numpy()
        adv, ret = ref_loader.reference_gae(r, v, d, lv, ld, 0.99, 0.95, use_gae=True)
        adv_mc, ret_mc = ref_loader.reference_gae(r, v, d, lv, ld, 0.99, 0.95, use_gae=False)
        save(f"gae_s{seed}", rewards=r, values=v, dones=d, last_value=lv, last_done=ld,
             gamma=0.99, gae_lambda=0.95, adv=adv, ret=ret, adv_mc=adv_mc, ret_mc=ret_mc)


def gen_gae32():
    # the legacy experiences path: both sibling sites (ppo.py / ippo.py), fractional next_done included
    for seed, (T, N, site) in [(0, (16, 8, "ppo")), (1, (128, 64, "ppo")), (2, (37, 33, "ippo")),
                               (3, (1, 5, "ippo")), (4, (64, 1, "ppo"))]:
        g = torch.Generator().manual_seed(100 + seed)
        r = torch.randn(T, N, generator=g).numpy()
        v = torch.randn(T, N, generator=g).numpy()
        d = (torch.rand(T, N, generator=g) < 0.05).numpy()
        nv = torch.randn(N, generator=g).numpy()
        nd = (torch.rand(N, generator=g) < 0.2).float().numpy()
        gamma, lam = (0.99, 0.95) if seed != 2 else (0.997, 0.9)
        adv, ret = ref_loader.reference_gae_f32(r, v, d, nv, nd, gamma, lam, site=site)
        save(f"gae32_s{seed}", rewards=r, values=v, dones=d, next_value=nv, next_done=nd,
             gamma=gamma, gae_lambda=lam, adv=adv, ret=ret)


def llm_case(seed, B, S, n_turns):
    """Synthetic multi-turn completions: per row a prompt, then up to n_turns generation turns separated
    by observation tokens, then padding; values zero outside the action mask, as ppo_llm.py:484 does."""
    g = torch.Generator().manual_seed(200 + seed)
    tid = np.full((B, S), -1, dtype=np.int64)
    for b in range(B):
        pos = int(torch.randint(1, max(2, S // 8), (1,), generator=g))
        k = 0
        turns = int(torch.randint(1, n_turns + 1, (1,), generator=g))
        while k < turns and pos < S - 1:
            ln = int(torch.randint(1, max(2, S // (2 * n_turns)), (1,), generator=g))
            tid[b, pos:min(S, pos + ln)] = k
            pos += ln + int(torch.randint(0, 4, (1,), generator=g))
            k += 1
    tid[0, :] = -1
    tid[0, S // 2] = n_turns - 1 if B > 1 else 0   # a row whose only turn is the last id
    mask = tid >= 0
    v = torch.randn(B, S, generator=g).numpy() * mask
    r = (torch.randn(B, S, generator=g).numpy() * 0.05 + (torch.rand(B, S, generator=g) < 0.05).numpy()) * mask
    return r.astype(np.float32), v.astype(np.float32), mask, tid


def gen_llm_gae():
    for seed, (B, S, K) in [(0, (4, 24, 2)), (1, (33, 260, 5)), (2, (7, 129, 3)), (3, (64, 1024, 6))]:
        r, v, mask, tid = llm_case(seed, B, S, K)
        out = {}
        for w in (False, True):
            ret, adv = ref_loader.reference_llm_gae(r, v, mask, 0.99, 0.95, w)
            out[f"tok_ret_w{int(w)}"], out[f"tok_adv_w{int(w)}"] = ret, adv
            for red in ("final_value", "mean"):
                ret, adv = ref_loader.reference_llm_gae(r, v, mask, 0.99, 0.95, w, turn_ids=tid,
                                                        turn_value_reduction=red)
                out[f"turn_ret_{red}_w{int(w)}"], out[f"turn_adv_{red}_w{int(w)}"] = ret, adv
        save(f"llm_gae_s{seed}", rewards=r, values=v, mask=mask, turn_ids=tid, gamma=0.99, gae_lambda=0.95, **out)


def gen_advnorm():
    for seed, shape in [(0, (7,)), (1, (128, 64)), (2, (131072,)), (3, (2,))]:
        g = torch.Generator().manual_seed(300 + seed)
        a = (torch.randn(*shape, generator=g) * 3 + 0.7).numpy()
        save(f"advnorm_s{seed}", advantages=a, normalized=ref_loader.reference_normalize_advantages(a))


def gen_per():
    ref = ref_loader.load_reference()
    TD = ref.TensorDict
    for seed, max_size, B in [(0, 100, 16), (1, 256, 64), (2, 1000, 128)]:
        torch.manual_seed(seed)
        g = torch.Generator().manual_seed(1000 + seed)
        buf = ref.PrioritizedReplayBuffer(max_size=max_size, alpha=0.6)
        rec = {}
        # fill beyond capacity in uneven batches so tree_ptr wraps at max_size
        n_added = 0
        step = 0
        sizes = []
        while n_added < max_size + max_size // 3:
            n = int(torch.randint(1, 17, (1,), generator=g))
            sizes.append(n)
            td = TD({"obs": torch.randn(n, 4, generator=g), "reward": torch.randn(n, generator=g)},
                    batch_size=[n])
            buf.add(td)
            n_added += n
            step += 1
            if step == 3:  # early update so max_priority grows before later adds
                idx = torch.randint(0, buf.size, (8,), generator=g)
                pr = (torch.randn(8, generator=g).abs() * 3 + 1e-6).numpy()
                buf.update_priorities(idx, pr)
                rec["early_idx"], rec["early_prio"], rec["early_at"] = idx.numpy(), pr, n_added
        rec["add_sizes"] = np.array(sizes)
        rec["tree_ptr"] = buf.tree_ptr
        rec["size"] = buf.size
        # updates with forced duplicates and a tiny priority (1e-5 clamp)
        idx = torch.randint(0, max_size, (B,), generator=g)
        idx[B // 2:] = idx[: B - B // 2]
        pr = (torch.randn(B, generator=g).abs() + 1e-6).numpy().astype(np.float32)
        pr[0] = 1e-9
        buf.update_priorities(idx, pr)
        rec["upd_idx"], rec["upd_prio"] = idx.numpy(), pr
        rec["sum_tree_after_update"] = np.array(buf.sum_tree.tree)
        rec["min_tree_after_update"] = np.array(buf.min_tree.tree)
        rec["max_priority"] = buf.max_priority
        # sampling with known uniforms: torch.rand(1) per stratum from the global generator
        torch.manual_seed(77 + seed)
        u = torch.rand(B)
        torch.manual_seed(77 + seed)
        indices = buf._sample_proportional(B)
        weights = buf._calculate_weights(indices, 0.4)
        rec["u"], rec["sample_idx"], rec["weights"], rec["beta"] = u.numpy(), indices.numpy(), weights.numpy(), 0.4
        save(f"per_s{seed}", max_size=max_size, alpha=0.6, **rec)


def gen_ring_nstep():
    ref = ref_loader.load_reference()
    TD = ref.TensorDict
    for seed, (max_size, N, n_step) in [(0, (10, 4, 3)), (1, (50, 16, 3)), (2, (23, 5, 5))]:
        g = torch.Generator().manual_seed(seed)
        steps = 3 * max_size // N + 7
        obs = torch.randint(0, 256, (steps, N, 6), generator=g, dtype=torch.uint8)
        nobs = torch.randint(0, 256, (steps, N, 6), generator=g, dtype=torch.uint8)
        act = torch.randint(0, 4, (steps, N), generator=g).float()
        rew = torch.randn(steps, N, generator=g)
        done = (torch.rand(steps, N, generator=g) < 0.08).float()
        ring = ref.ReplayBuffer(max_size=max_size)
        nbuf = ref.MultiStepReplayBuffer(max_size=max_size, n_step=n_step, gamma=0.99)
        returned_none = []
        for s in range(steps):
            td = TD({"obs": obs[s], "action": act[s], "reward": rew[s], "next_obs": nobs[s],
                     "done": done[s]}, batch_size=[N])
            ring.add(td.clone())
            out = nbuf.add(td.clone())
            returned_none.append(out is None)
        save(f"ring_nstep_s{seed}", max_size=max_size, N=N, n_step=n_step, gamma=0.99,
             obs=obs.numpy(), next_obs=nobs.numpy(), action=act.numpy(), reward=rew.numpy(),
             done=done.numpy(), returned_none=np.array(returned_none),
             ring_cursor=ring._cursor, ring_size=ring.size,
             **{f"ring_{k}": v.numpy() for k, v in ring.storage.items()},
             nstep_cursor=nbuf._cursor, nstep_size=nbuf.size,
             **{f"nstep_{k}": v.numpy() for k, v in nbuf.storage.items()})


def gen_rng():
    rec = {}
    cases = [(0, 100000, 64), (7, 1000000, 1024), (3, 50, 50), (5, 1000, 999), (11, 4097, 2048),
             (2, 10_000_000, 1024), (13, 1, 1), (17, 2, 2)]
    for k, (seed, n, B) in enumerate(cases):
        torch.manual_seed(seed)
        rec[f"perm_{k}"] = torch.randperm(n)[:B].numpy()
    rec["cases"] = np.array(cases)
    for seed in (0, 1, 123456789):
        torch.manual_seed(seed)
        rec[f"rand_{seed}"] = torch.rand(700).numpy()  # crosses the 624-word twist boundary
    save("torch_rng", **rec)


def gen_bptt():
    """Sequence cutting + padding by the extracted reference functions (ref_loader.reference_bptt)."""
    cases = [(0, 12, 3, None, "chunked", 0.2), (1, 12, 3, 3, "chunked", 0.2), (2, 12, 3, 3, "maximum", 0.2),
             (3, 16, 4, 4, "fifty_percent_overlap", 0.15), (4, 40, 7, 5, "chunked", 0.05),
             (5, 40, 7, 8, "maximum", 0.0), (6, 33, 5, 1, "chunked", 0.3), (7, 9, 2, 20, "chunked", 0.1),
             (8, 64, 33, 6, "fifty_percent_overlap", 0.08)]
    for seed, T, N, msl, tp, pd in cases:
        g = torch.Generator().manual_seed(100 + seed)
        d = (torch.rand(T, N, generator=g) < pd).numpy()
        x = torch.randn(T, N, 3, generator=g).numpy()
        slices, padded, mask, L, S = ref_loader.reference_bptt(d, x, msl, tp)
        save(f"bptt_s{seed}", dones=d, data=x, max_seq_len=np.array(-1 if msl is None else msl),
             seq_type=np.array(tp), unpadded=np.concatenate(slices),
             lengths=np.array([len(s) for s in slices]), padded=padded, pad_mask=mask,
             max_sequence_length=np.array(L), num_sequences=np.array(S))


def marb_scenario(seed: int):
    """Deterministic multi-agent scenario: (memory_size, fields, agents, list of save calls,
    list of sample batch sizes).  Shared by the golden generator and the GPU parity test."""
    rng = np.random.default_rng(1000 + seed)
    agents = ["agent_0", "agent_1", "other"]
    fields = ["state", "action", "reward", "next_state", "done"]
    cap = [16, 40, 9][seed]
    vect = seed == 1
    n_env = 4
    saves = []
    for step in range([25, 14, 30][seed]):
        def obs(a):
            lead = (n_env,) if vect else ()
            if a == "other" and seed != 1:
                return {"img": rng.integers(0, 256, (*lead, 2, 3, 3)).astype(np.uint8),
                        "vec": rng.standard_normal((*lead, 5)).astype(np.float32)}
            if a == "agent_1" and seed == 2:
                return (rng.standard_normal((*lead, 2)), rng.integers(0, 9, (*lead, 1)))
            return rng.standard_normal((*lead, 4)).astype(np.float32)
        state = {a: obs(a) for a in agents}
        nstate = {a: obs(a) for a in agents}
        if vect:
            action = {a: rng.integers(0, 5, (n_env,)) for a in agents}
            reward = {a: rng.standard_normal(n_env) for a in agents}
            done = {a: rng.random(n_env) < 0.2 for a in agents}
        else:
            action = {a: int(rng.integers(0, 5)) if a != "other" else rng.standard_normal(2) for a in agents}
            reward = {a: float(rng.standard_normal()) for a in agents}
            done = {a: bool(rng.random() < 0.2) for a in agents}
        saves.append(((state, action, reward, nstate, done), vect))
    return cap, fields, agents, saves, [3, 7, 1, 8]


def flatten_marb(sample, fields, agents):
    out = {}
    for f, per_agent in zip(fields, sample):
        for a in agents:
            v = per_agent[a]
            if isinstance(v, dict):
                for k, t in v.items():
                    out[f"{f}|{a}|d:{k}"] = t
            elif isinstance(v, tuple):
                for i, t in enumerate(v):
                    out[f"{f}|{a}|t:{i}"] = t
            else:
                out[f"{f}|{a}"] = v
    return out


def gen_marb():
    """MultiAgentReplayBuffer: the unmodified reference class through save_to_memory / sample."""
    import random

    M = ref_loader.load_reference_marb()
    for seed in range(3):
        cap, fields, agents, saves, batches = marb_scenario(seed)
        buf = M(cap, fields, agents, device="cpu")
        arrays = {}
        random.seed(500 + seed)
        k = 0
        for i, (args, vect) in enumerate(saves):
            buf.save_to_memory(*args, is_vectorised=vect)
            if i in (len(saves) // 3, len(saves) - 1):
                for bs in batches:
                    if bs <= len(buf):
                        s = flatten_marb(buf.sample(bs), fields, agents)
                        for name, t in s.items():
                            assert t.dtype == torch.float32
                            arrays[f"s{k}|{name}"] = t.numpy()
                        k += 1
        arrays["n_samples"] = np.array(k)
        arrays["final_len"] = np.array(len(buf))
        arrays["counter"] = np.array(buf.counter)
        save(f"marb_s{seed}", **arrays)


if __name__ == "__main__":
    if sys.argv[1:] == ["marb"]:
        gen_marb()
        sys.exit(0)
    if sys.argv[1:] == ["gae32"]:
        gen_gae32()
        sys.exit(0)
    if sys.argv[1:] == ["advnorm"]:
        gen_advnorm()
        sys.exit(0)
    if sys.argv[1:] == ["llm_gae"]:
        gen_llm_gae()
        sys.exit(0)
    if sys.argv[1:] == ["bptt"]:
        gen_bptt()
        sys.exit(0)
    gen_gae()
    gen_gae32()
    gen_llm_gae()
    gen_advnorm()
    gen_per()
    gen_ring_nstep()
    gen_rng()
    gen_bptt()
    gen_marb()
