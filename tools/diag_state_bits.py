"""Scratch diagnostic: bitwise state differences (device vs compiled reference) in the 2400-arena 3v3 case."""
import os, sys, ctypes as C, struct
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import importlib
ggl = importlib.import_module("reinforcement-learning_b200")
import env_parity, oracle_ref
import torch

TARGET, CAR = 337, 2
num_arenas, ppt, seed = 2400, 3, 43
P = 2 * ppt

def flat(x):
    if isinstance(x, C.Array):
        return [v for e in x for v in flat(e)]
    if isinstance(x, C.Structure):
        return [v for f in x._fields_ for v in flat(getattr(x, f[0]))]
    return [x]

def bits(v):
    return struct.unpack("<I", struct.pack("<f", v))[0] if isinstance(v, float) else int(v)

L = oracle_ref.lib(); L.ref_set_threads(1); L.ref_seed(99 + seed)
ref = oracle_ref.RefEnvSet(num_arenas, ppt, obs_kind=1, state_setter=oracle_ref.abi.SETTER_INJECTED)
dev = ggl.EnvSet(num_arenas, ppt, obs_kind=1, state_setter=oracle_ref.abi.SETTER_INJECTED, seed=seed)
ids, teams = ref.slot_info(); dev.set_slot_order(ids)
for a in range(num_arenas):
    st = ref.get_state(a); dev.set_state(a, st); ref.set_reset_state(a, st); dev.set_reset_state(a, st)
rng = np.random.default_rng(seed)
acts_d = torch.zeros(dev.num_players, dtype=torch.int32, device=dev.device)
obs_r, masks_r, _, _ = ref.outputs()
for s in range(8):
    acts = env_parity.random_valid_actions(rng, masks_r)
    ref.step(acts); acts_d.copy_(torch.from_numpy(acts)); dev.step(acts_d); dev.sync()
    obs_r, masks_r, rew_r, term_r = ref.outputs()
    # bitwise census over a sample of arenas
    nd = nz = nt = 0
    for a in list(range(0, num_arenas, 40)) + [TARGET]:
        sr, sd = ref.get_state(a), dev.get_state(a)
        for c in range(sr.num_cars):
            cr = sr.cars[c]; cd = [sd.cars[k] for k in range(sd.num_cars) if sd.cars[k].id == cr.id][0]
            for n in ("pos", "rot", "vel", "ang_vel"):
                for x, y in zip(flat(getattr(cd, n)), flat(getattr(cr, n))):
                    nt += 1
                    if bits(x) != bits(y):
                        if x == y: nz += 1
                        else: nd += 1
    print("step", s, "floats compared", nt, "differ in value", nd, "differ in sign of zero only", nz)
    if s in (4, 5, 6):
        sr, sd = ref.get_state(TARGET), dev.get_state(TARGET)
        cr = [sr.cars[k] for k in range(sr.num_cars) if sr.cars[k].id == CAR][0]
        cd = [sd.cars[k] for k in range(sd.num_cars) if sd.cars[k].id == CAR][0]
        for n in ("pos", "rot", "vel", "ang_vel"):
            print("   car2", n, "dev", [repr(v) for v in flat(getattr(cd, n))], "\n        ", " " * len(n), "ref", [repr(v) for v in flat(getattr(cr, n))])
        print("   contact flags dev", cd.world_contact_has, list(cd.world_contact_normal), "ref", cr.world_contact_has, list(cr.world_contact_normal))
ref.close(); dev.close()
