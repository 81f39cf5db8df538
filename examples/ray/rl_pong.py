"""Policy-gradient Pong with parallel roll-out actors (reference example: pyzoo/zoo/examples/ray/rl_pong/rl_pong.py -- `batch_size`
PongEnv actors each play one episode with the current policy and return its gradient; the driver averages the gradients and takes an
RMSProp step).  The actors run on RayContext's worker pool.

No gym / Atari ROM in this image, so the environment is a small Pong written here: a ball bounces in a 16 x 16 court, the agent moves
its paddle up or down, an episode is `POINTS` serves and a point is +1 (returned) or -1 (missed).  The observation is the
difference of two rendered frames, as in the reference; the policy is the same two-layer network trained with the same update."""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np

from zoo.ray import RayContext

H, SIZE, GAMMA, DECAY, LR = 64, 16, 0.95, 0.99, 1e-2
D = SIZE * SIZE


class MiniPong:
    """Ball from the left wall towards the agent's paddle (right wall, 3 cells tall)."""

    def __init__(self, seed, points):
        self.rng, self.points = np.random.RandomState(seed), points

    def _serve(self):
        self.bx, self.by = 0.0, float(self.rng.randint(2, SIZE - 2))
        self.vx, self.vy = 1.0, float(self.rng.choice([-1.0, -0.5, 0.5, 1.0]))

    def reset(self):
        self.left, self.paddle = self.points, SIZE // 2
        self._serve()
        return self._frame()

    def _frame(self):
        f = np.zeros((SIZE, SIZE), np.float32)
        f[int(round(self.by)) % SIZE, min(int(self.bx), SIZE - 1)] = 1.0
        f[max(self.paddle - 1, 0):self.paddle + 2, SIZE - 1] = 1.0
        return f.ravel()

    def step(self, up):
        self.paddle = int(np.clip(self.paddle + (-1 if up else 1), 1, SIZE - 2))
        self.bx += self.vx
        self.by += self.vy
        if self.by < 0 or self.by > SIZE - 1:
            self.vy, self.by = -self.vy, float(np.clip(self.by, 0, SIZE - 1))
        reward = 0.0
        if self.bx >= SIZE - 1:
            reward = 1.0 if abs(self.by - self.paddle) <= 1.5 else -1.0
            self.left -= 1
            self._serve()
        return self._frame(), reward, self.left == 0


def discount_rewards(r):
    out, run = np.zeros_like(r), 0.0
    for t in reversed(range(r.size)):
        if r[t] != 0:
            run = 0.0                    # a point ends the rally
        run = run * GAMMA + r[t]
        out[t] = run
    return out


def policy_forward(x, model):
    h = np.maximum(model["W1"] @ x, 0.0)
    return 1.0 / (1.0 + np.exp(-(model["W2"] @ h))), h


def policy_backward(eph, epx, epdlogp, model):
    dW2 = (eph.T @ epdlogp).ravel()
    dh = np.outer(epdlogp, model["W2"])
    dh[eph <= 0] = 0
    return {"W1": dh.T @ epx, "W2": dW2}


class PongEnv:
    def __init__(self, seed, points):
        self.env, self.rng = MiniPong(seed, points), np.random.RandomState(1000 + seed)

    def compute_gradient(self, model):
        obs, prev = self.env.reset(), None
        xs, hs, dlogps, drs, total, done = [], [], [], [], 0.0, False
        while not done:
            x = obs - prev if prev is not None else np.zeros(D, np.float32)
            prev = obs
            aprob, h = policy_forward(x, model)
            up = self.rng.uniform() < aprob
            xs.append(x), hs.append(h), dlogps.append((1.0 if up else 0.0) - aprob)
            obs, reward, done = self.env.step(up)
            total += reward
            drs.append(reward)
        adv = discount_rewards(np.asarray(drs, np.float32))
        adv = (adv - adv.mean()) / (adv.std() + 1e-8)
        return policy_backward(np.vstack(hs), np.vstack(xs), np.asarray(dlogps, np.float32) * adv, model), total


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch_size", type=int, default=4, help="roll-out actors per model update")
    ap.add_argument("--iterations", type=int, default=int(os.environ.get("PONG_ITERS", 400)))
    ap.add_argument("--points", type=int, default=10, help="serves per episode")
    args = ap.parse_args()
    rng = np.random.RandomState(0)
    model = {"W1": rng.randn(H, D).astype(np.float32) / np.sqrt(D), "W2": rng.randn(H).astype(np.float32) / np.sqrt(H)}
    cache = {k: np.zeros_like(v) for k, v in model.items()}
    ctx = RayContext(sc=None, object_store_memory="1g").init()
    actors = [ctx.actor(PongEnv, seed, args.points) for seed in range(args.batch_size)]
    running, first = None, None
    for it in range(1, args.iterations + 1):
        results = RayContext.get_objects([a.compute_gradient.remote(model) for a in actors])
        grad = {k: np.mean([g[k] for g, _ in results], axis=0) for k in model}
        reward = float(np.mean([r for _, r in results]))
        running = reward if running is None else 0.9 * running + 0.1 * reward
        first = running if it == 10 else first
        for k in model:                                             # RMSProp ascent on the expected reward
            cache[k] = DECAY * cache[k] + (1 - DECAY) * grad[k] ** 2
            model[k] += LR * grad[k] / (np.sqrt(cache[k]) + 1e-5)
        if it % 50 == 0:
            print("iteration %d: batch mean reward %.2f, running mean %.2f" % (it, reward, running))
    print("running mean reward: %.2f after 10 updates -> %.2f after %d (episode = %d points, range -%d..%d)" %
          (first if first is not None else running, running, args.iterations, args.points, args.points, args.points))
    ctx.stop()
