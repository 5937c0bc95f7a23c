"""Randomised parity of ``populate_rl_data`` (host scalars + K4 on the GPU) against the REFERENCE's own pandas implementation
(``pipelinerl.finetune.rl.populate_rl_data`` from ``oracle/_ref``, rl/__init__.py:380-501; the oracle restatement when
``oracle/_ref`` is not staged).  TEST INFRASTRUCTURE.  Every output list is compared with ``==`` on the Python floats: bit-exact f64.

    python tests/fuzz_advantage.py <first_seed> <n_seeds>      (GPU box)

Draws: group sizes 1..17 (single-member groups: pandas std is NaN -> nan_to_num), equal rewards inside a group (std 0), binary /
fractional / tiny / huge / negative rewards, several ``step_index`` values per ``group_id`` (the key is ``group_id + "_" + step_index``,
:395), group ids that contain "_" and look like other keys ("g1" step 10 vs "g1_1" step 0), rows shuffled so that a group's members
are not contiguous, rollouts with and without the EOS token (overflow), prompts of any length (num_labels 0..n), per-token rewards that
differ from the first token's, and ``divide_advantage_by_std`` on / off.
"""
from __future__ import annotations

import copy
import os
import random
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

KEYS = ("advantages", "group_tokens", "overflow", "num_labels")


def _reference():
    try:
        from oracle import build_ref
        rl = build_ref.load()
        return "reference", rl.populate_rl_data, rl.RLConfig
    except Exception:  # noqa: BLE001
        import oracle
        return "port", oracle.oracle_populate_rl_data, oracle.OracleRLConfig


def draw_dataset(rnd: random.Random):
    eos = 2
    reward_kind = rnd.choice(["binary", "fraction", "tiny", "huge", "mixed"])

    def reward():
        if reward_kind == "binary":
            return float(rnd.randint(0, 1))
        if reward_kind == "fraction":
            return rnd.choice([0.1, 1.0 / 3.0, 0.7, -0.25, 1.0])
        if reward_kind == "tiny":
            return rnd.random() * 1e-6
        if reward_kind == "huge":
            return rnd.choice([1e6, 1e6 + 1.0, -3e5, 12345.678])
        return rnd.uniform(-2.0, 2.0)

    rows = []
    gids = [rnd.choice(["g", "task_7", "a_b_c", "1"]) + str(i) for i in range(rnd.randint(1, 12))]
    if rnd.random() < 0.5:
        gids += ["g1_1", "g1"]                       # "g1" + step 10 vs "g1_1" + step 0 share the prefix "g1_1"
    for gid in dict.fromkeys(gids):                  # unique (group, step, rollout) triples, as the reference asserts (:405)
        for step in rnd.sample([0, 1, 2, 10], rnd.randint(1, 2)):
            same = rnd.random() < 0.25
            r0 = reward()
            for a in range(rnd.choice([1, 1, 2, 3, 4, 8, 16, 17])):
                n = rnd.randint(1, 60)
                k = rnd.randint(0, n)
                ids = [rnd.randint(3, 1000) for _ in range(n)]
                if rnd.random() < 0.7:
                    ids[rnd.randrange(n)] = eos
                r = r0 if same else reward()
                rewards = [r] * n
                if rnd.random() < 0.15 and n > 1:    # the reference uses rewards[0] for the statistics and every token's own reward for its advantage
                    rewards[-1] = r + 0.5
                rows.append({"group_id": gid, "rollout_index": a, "step_index": step, "input_ids": ids, "labels": [-100] * k + ids[k:],
                             "rewards": rewards, "advantages": [0.0] * n, "overflow": [0] * n, "group_tokens": [0] * n, "num_labels": [0] * n,
                             "old_logprobs": [0.0] * n, "ref_logprobs": [0.0] * n})
    if rnd.random() < 0.6:
        rnd.shuffle(rows)
    return rows, eos, dict(divide_advantage_by_std=rnd.random() < 0.7)


def fuzz(seed: int, n_cases: int = 8):
    from pipelinerl_swe_b200 import RLConfig, populate_rl_data
    kind, ref_fn, RefCfg = _reference()
    rnd = random.Random(seed)
    bad = []
    for c in range(n_cases):
        rows, eos, cfg = draw_dataset(rnd)
        want = ref_fn(copy.deepcopy(rows), eos, RefCfg(**cfg))
        got = populate_rl_data(rows, eos, RLConfig(**cfg))
        assert got is rows                                           # mutated in place and returned, like the reference
        for i, (a, b) in enumerate(zip(got, want)):
            for k in KEYS:
                # values with ==; kinds: int where the reference has an int, float where it has a float (np.float64 is a float subclass)
                if a[k] != b[k] or any(isinstance(x, int) != isinstance(y, int) for x, y in zip(a[k][:1], b[k][:1])):
                    bad.append((c, i, k, a[k][:2], b[k][:2], a["group_id"], cfg))
                    break
    return not bad, f"seed {seed}: {n_cases} datasets vs {kind} ({'pandas' if kind == 'reference' else 'restatement'}), {len(bad)} differing rows {bad[:2]}"


if __name__ == "__main__":
    first, n = int(sys.argv[1]), int(sys.argv[2])
    n_bad = 0
    for s in range(first, first + n):
        ok, msg = fuzz(s)
        n_bad += not ok
        print("FUZZ", "OK" if ok else "FAIL", msg, flush=True)
    print(f"FUZZ SUMMARY seeds {first}..{first + n - 1}: {n_bad} failing seeds")
    sys.exit(1 if n_bad else 0)
