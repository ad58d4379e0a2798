"""TEST INFRASTRUCTURE (oracle) — numpy restatement of GGL::BatchedWelfordStat
(GigaLearnCPP/src/private/GigaLearnCPP/Util/WelfordStat.h:69-243): IncrementRow (:89-97), GetSTD (:107-127), NormalizeInPlace
(:131-216, coefficient formulas :139-160, `row[j] * invSTD[j] + negMeanDivSTD[j]` as separate float32 multiply and add).
Pinned bit-exact against the compiled reference (oracle/_ref/libref_welford.so, tests/test_obsstat_oracle_cpu.py) and against
tests/golden/obsstat_golden.npz. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline may import it."""
import numpy as np


class BatchedWelfordStat:
    def __init__(self, width):
        self.width = width
        self.means = np.zeros(width, np.float64)
        self.vars = np.zeros(width, np.float64)
        self.count = 0

    def increment_row(self, row):                              # :89-97 (vectorised over the independent columns)
        delta = row.astype(np.float64) - self.means
        delta_n = delta / np.float64(self.count + 1)
        self.means = self.means + delta_n
        self.vars = self.vars + delta * delta_n * np.float64(self.count)
        self.count += 1

    def increment_rows(self, obs, idx):
        for i in idx:
            self.increment_row(obs[int(i)])

    def get_std(self):                                         # :107-127
        if self.count < 2:
            return np.ones(self.width)
        var = self.vars / np.float64(self.count - 1)
        return np.where(var > 0, np.sqrt(np.where(var > 0, var, 1.0)), 1.0)

    def normalize_in_place(self, obs, mean_clamp, min_std):    # :131-216
        if self.count < 2 or obs.shape[0] <= 0:
            return
        cm = np.clip(self.means, -mean_clamp, mean_clamp)
        cs = np.maximum(self.get_std(), min_std)
        inv = (np.float32(1.0) / cs.astype(np.float32)).astype(np.float32)
        neg = (-(cm.astype(np.float32)) * inv).astype(np.float32)
        obs[...] = (obs * inv[None, :]).astype(np.float32) + neg[None, :]
