"""Divergence rates of the loaded build against the compiled reference in steady-state play (tests/env_parity.run_steady_state):
arena-steps whose obs leave the 1e-4 tolerance although the physics is re-synchronised before every step.
    GGL_LIB_PATH=... python tools/parity_rates.py [arenas] [steps]"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import env_parity   # noqa: E402
import scenarios    # noqa: E402

ggl = importlib.import_module("reinforcement-learning_b200")
arenas = int(sys.argv[1]) if len(sys.argv) > 1 else 256
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 60
for ppt, scripted in ((2, False), (2, True), (3, True), (1, True)):
    n = arenas if ppt < 3 else arenas // 2
    kw = dict(scripts=scenarios.env_scripts(n, ppt, 2)) if scripted else {}
    r = env_parity.run_steady_state(ggl, num_arenas=n, ppt=ppt, preroll=160 if not scripted else 100, steps=steps, seed=21 + ppt, **kw)
    print("%dv%d %-10s arena_steps %6d bad %4d (%.2f %%)  obs_bad %5d  mask/term/event mismatches %d/%d/%d  int_mismatch_on_bad %d  rew %.1e  events %s" %
          (ppt, ppt, "scripted" if scripted else "unscripted", r["arena_steps"], r["bad_arena_steps"], 100.0 * r["bad_arena_steps"] / max(1, r["arena_steps"]),
           r["obs_bad"], r["mask_mismatch"], r["term_mismatch"], r["event_mismatch"], r["int_mismatch_on_bad_steps"], r["rew"], list(map(int, r["events"]))), flush=True)
