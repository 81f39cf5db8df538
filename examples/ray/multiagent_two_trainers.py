"""Two training methods at once in one multi-agent environment (reference example:
pyzoo/zoo/examples/ray/rllib/multiagent_two_trainers.py -- four independent CartPole agents, even agents driven by a PPO policy,
odd agents by a DQN policy, each trainer improves its own policy and the weights are synced to the other after every round).

RLlib is not in this image, so the two trainers are written here in torch at example scale: a clipped-surrogate policy-gradient
trainer ("ppo_policy") and a double-DQN trainer with a replay buffer ("dqn_policy").  Each is an actor on RayContext's worker pool;
the driver moves the policies' weights between them exactly where the reference calls `set_weights(get_weights([...]))`."""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import torch
import torch.nn as nn

from zoo.ray import RayContext


class CartPole:
    """The classic cart-pole balance task (Barto, Sutton & Anderson), Euler-integrated; episode ends at |x| > 2.4, |theta| > 12 deg
    or 200 steps; reward 1 per step."""

    def __init__(self, seed):
        self.rng = np.random.RandomState(seed)

    def reset(self):
        self.s, self.t = self.rng.uniform(-0.05, 0.05, 4), 0
        return self.s.astype(np.float32)

    def step(self, action):
        x, xd, th, thd = self.s
        f = 10.0 if action == 1 else -10.0
        ct, st = np.cos(th), np.sin(th)
        tmp = (f + 0.05 * thd * thd * st) / 1.1
        tha = (9.8 * st - ct * tmp) / (0.5 * (4.0 / 3.0 - 0.1 * ct * ct / 1.1))
        xa = tmp - 0.05 * tha * ct / 1.1
        self.s = np.array([x + 0.02 * xd, xd + 0.02 * xa, th + 0.02 * thd, thd + 0.02 * tha])
        self.t += 1
        done = abs(self.s[0]) > 2.4 or abs(self.s[2]) > 0.2095 or self.t >= 200
        return self.s.astype(np.float32), 1.0, done


class MultiCartpole:
    """`n` independent cart-poles stepped together: dicts keyed by agent id in, dicts out; finished agents restart."""

    def __init__(self, n, seed=0):
        self.envs = [CartPole(seed * 100 + i) for i in range(n)]

    def reset(self):
        return {i: e.reset() for i, e in enumerate(self.envs)}

    def step(self, actions):
        obs, rew, done = {}, {}, {}
        for i, a in actions.items():
            o, r, d = self.envs[i].step(a)
            obs[i], rew[i], done[i] = (self.envs[i].reset() if d else o), r, d
        return obs, rew, done


def policy_mapping_fn(agent_id):
    return "ppo_policy" if agent_id % 2 == 0 else "dqn_policy"


def make_policies():
    torch.manual_seed(0)
    return {"ppo_policy": nn.Sequential(nn.Linear(4, 64), nn.Tanh(), nn.Linear(64, 2)),          # action logits
            "dqn_policy": nn.Sequential(nn.Linear(4, 64), nn.ReLU(), nn.Linear(64, 2))}          # action values


class Trainer:
    """Holds BOTH policies, acts with both in the shared environment, trains only `policy_to_train`."""

    def __init__(self, policy_to_train, seed):
        torch.manual_seed(seed)
        torch.set_num_threads(1)
        self.policies, self.train_id = make_policies(), policy_to_train
        self.env, self.rng = MultiCartpole(4, seed), np.random.RandomState(seed)
        self.opt = torch.optim.Adam(self.policies[policy_to_train].parameters(), lr=3e-3 if policy_to_train == "ppo_policy" else 1e-3)
        self.target = make_policies()["dqn_policy"]
        self.replay, self.eps, self.steps = [], 1.0, 0

    def get_weights(self, ids):
        return {i: {k: v.clone() for k, v in self.policies[i].state_dict().items()} for i in ids}

    def set_weights(self, weights):
        for i, sd in weights.items():
            self.policies[i].load_state_dict(sd)
        return True

    def _act(self, pid, obs, explore):
        with torch.no_grad():
            out = self.policies[pid](torch.from_numpy(obs))
        if pid == "ppo_policy":
            p = torch.softmax(out, -1).numpy()
            a = int(self.rng.rand() < p[1]) if explore else int(p[1] > 0.5)
            return a, float(np.log(p[a] + 1e-8))
        if explore and self.rng.rand() < self.eps:
            return int(self.rng.randint(2)), 0.0
        return int(out.argmax()), 0.0

    def train(self, env_steps=800):
        obs = self.env.reset()
        traj = {i: [] for i in obs}                      # per agent: (obs, action, logp, reward, next_obs, done)
        returns, ep = {"ppo_policy": [], "dqn_policy": []}, {i: 0.0 for i in obs}
        for _ in range(env_steps):
            acts = {}
            for i, o in obs.items():
                pid = policy_mapping_fn(i)
                acts[i] = self._act(pid, o, explore=(pid == self.train_id))          # the other trainer's policy acts greedily
            nobs, rew, done = self.env.step({i: a for i, (a, _) in acts.items()})
            for i in obs:
                traj[i].append((obs[i], acts[i][0], acts[i][1], rew[i], nobs[i], done[i]))
                ep[i] += rew[i]
                if done[i]:
                    returns[policy_mapping_fn(i)].append(ep[i])
                    ep[i] = 0.0
            obs = nobs
        mine = [t for i, ts in traj.items() if policy_mapping_fn(i) == self.train_id for t in [ts]]
        (self._train_ppo if self.train_id == "ppo_policy" else self._train_dqn)(mine)
        return {"policy_reward_mean": {k: (float(np.mean(v)) if v else float("nan")) for k, v in returns.items()}}

    def _train_ppo(self, trajectories):
        o, a, lp, adv = [], [], [], []
        for ts in trajectories:
            g, gs = 0.0, []
            for (_, _, _, r, _, d) in reversed(ts):
                g = r + 0.98 * (0.0 if d else g)
                gs.append(g)
            gs.reverse()
            o += [t[0] for t in ts]; a += [t[1] for t in ts]; lp += [t[2] for t in ts]; adv += gs
        o, a = torch.from_numpy(np.stack(o)), torch.tensor(a)
        lp, adv = torch.tensor(lp, dtype=torch.float32), torch.tensor(adv, dtype=torch.float32)
        adv = (adv - adv.mean()) / (adv.std() + 1e-6)
        for _ in range(6):                                # several epochs on the batch, ratio clipped to [0.8, 1.2]
            new = torch.log_softmax(self.policies["ppo_policy"](o), -1).gather(1, a[:, None]).squeeze(1)
            ratio = torch.exp(new - lp)
            loss = -torch.min(ratio * adv, ratio.clamp(0.8, 1.2) * adv).mean()
            self.opt.zero_grad(); loss.backward(); self.opt.step()

    def _train_dqn(self, trajectories):
        for ts in trajectories:
            self.replay += [(t[0], t[1], t[3], t[4], t[5]) for t in ts]
        self.replay = self.replay[-20000:]
        q = self.policies["dqn_policy"]
        for _ in range(200):
            idx = self.rng.randint(len(self.replay), size=64)
            o, a, r, no, d = (np.stack([self.replay[i][k] for i in idx]) for k in range(5))
            o, no, a = torch.from_numpy(o), torch.from_numpy(no), torch.from_numpy(a.astype(np.int64))
            with torch.no_grad():                         # double DQN target, gamma 0.95 as in the reference config
                best = q(no).argmax(-1, keepdim=True)
                tgt = torch.from_numpy(r.astype(np.float32)) + 0.95 * (1 - torch.from_numpy(d.astype(np.float32))) * self.target(no).gather(1, best).squeeze(1)
            loss = nn.functional.smooth_l1_loss(q(o).gather(1, a[:, None]).squeeze(1), tgt)
            self.opt.zero_grad(); loss.backward(); self.opt.step()
            self.steps += 1
            if self.steps % 100 == 0:
                self.target.load_state_dict(q.state_dict())
        self.eps = max(0.05, self.eps * 0.8)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--iterations", type=int, default=int(os.environ.get("RL_ITERS", 12)))
    args = ap.parse_args()
    ctx = RayContext(sc=None, object_store_memory="1g").init()
    ppo_trainer, dqn_trainer = ctx.actor(Trainer, "ppo_policy", 1), ctx.actor(Trainer, "dqn_policy", 2)
    first = last = None
    for i in range(args.iterations):
        d = dqn_trainer.train.remote()                    # the two trainers improve their policies side by side
        p = ppo_trainer.train.remote()
        d, p = d.get(), p.get()
        # swap weights to synchronize (reference: ppo_trainer.set_weights(dqn_trainer.get_weights(["dqn_policy"])) and vice versa)
        ppo_trainer.set_weights.remote(dqn_trainer.get_weights.remote(["dqn_policy"]).get()).get()
        dqn_trainer.set_weights.remote(ppo_trainer.get_weights.remote(["ppo_policy"]).get()).get()
        last = {"dqn_policy": d["policy_reward_mean"]["dqn_policy"], "ppo_policy": p["policy_reward_mean"]["ppo_policy"]}
        first = first or last
        print("== Iteration %d ==  policy_reward_mean: dqn_policy %.1f  ppo_policy %.1f" % (i, last["dqn_policy"], last["ppo_policy"]))
    print("episode reward (max 200): dqn_policy %.1f -> %.1f, ppo_policy %.1f -> %.1f" %
          (first["dqn_policy"], last["dqn_policy"], first["ppo_policy"], last["ppo_policy"]))
    ctx.stop()
