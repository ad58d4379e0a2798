"""TEST INFRASTRUCTURE (oracle) — CPU restatement of the trajectory assembly inside GGL::Learner::Start
(GigaLearnCPP/src/public/GigaLearnCPP/Learner.cpp): the `Trajectory` struct (:504-547), the obs/mask append before the env
step (:701-712), the result append, terminal decision, forced truncation and `combinedTraj.Append` after it (:823-863), the
per-iteration `combinedTraj.Clear()` (:631) and the stop rule `combinedTraj.Length() < tsPerItr` (:665).

PARITY UNPINNED: this logic lives in the middle of Learner::Start, which cannot run without the whole learner (policy networks,
metrics sender, render); the reference has no test or fixture for it. The restatement follows the cited lines one to one with
Python lists standing in for the std::vectors. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline may import it.
"""
import numpy as np

TRUNCATED = 2       # RLGC::TerminalType (RLGymCPP/src/RLGymCPP/TerminalConditions/TerminalCondition.h)


class _Trajectory:                                  # Learner.cpp:504-547
    def __init__(self):
        self.clear()

    def clear(self):                                # :524-532
        self.states, self.next_states, self.rewards, self.log_probs = [], [], [], []
        self.action_masks, self.terminals, self.actions = [], [], []

    def append(self, o):                            # :534-542
        self.states += o.states; self.next_states += o.next_states; self.rewards += o.rewards; self.log_probs += o.log_probs
        self.action_masks += o.action_masks; self.terminals += o.terminals; self.actions += o.actions

    def length(self):                               # :544-546
        return len(self.actions)


class TrajOracle:
    def __init__(self, num_players, players_per_arena, max_episode_length):
        self.n, self.ppa, self.max_len = num_players, players_per_arena, max_episode_length
        self.trajectories = [_Trajectory() for _ in range(num_players)]      # :549
        self.combined = _Trajectory()

    def begin_iteration(self):                      # :631
        self.combined.clear()

    def push_obs(self, obs, masks):                 # :701-712
        for p in range(self.n):
            t = self.trajectories[p]
            t.states.append(np.array(obs[p], np.float32))
            t.action_masks.append(np.array(masks[p], np.uint8))

    def push_result(self, actions, log_probs, rewards, terminals, obs_final):
        for p in range(self.n):                     # :823-829
            t = self.trajectories[p]
            t.actions.append(int(actions[p])); t.rewards.append(np.float32(rewards[p])); t.log_probs.append(np.float32(log_probs[p]))
        cur = np.zeros(self.n, np.int8)             # :831-841
        for a in range(self.n // self.ppa):
            if terminals[a]:
                cur[a * self.ppa:(a + 1) * self.ppa] = terminals[a]
        for p in range(self.n):                     # :844-862
            term = int(cur[p])
            t = self.trajectories[p]
            if not term and t.length() >= self.max_len:
                term = TRUNCATED
            t.terminals.append(term)
            if term:
                if term == TRUNCATED:
                    t.next_states.append(np.array(obs_final[p], np.float32))
                self.combined.append(t)
                t.clear()

    def length(self):
        return self.combined.length()

    def player_lengths(self):
        return np.array([t.length() for t in self.trajectories], np.int32)

    def tensors(self, obs_size, num_actions):
        c = self.combined
        L = c.length()
        return {"states": np.array(c.states, np.float32).reshape(L, obs_size), "action_masks": np.array(c.action_masks, np.uint8).reshape(L, num_actions),
                "actions": np.array(c.actions, np.int32), "log_probs": np.array(c.log_probs, np.float32),
                "rewards": np.array(c.rewards, np.float32), "terminals": np.array(c.terminals, np.int8),
                "next_states": np.array(c.next_states, np.float32).reshape(len(c.next_states), obs_size)}
