"""Generates tests/golden/*.pt by running the UNMODIFIED reference (imported from /root/reference).

TEST INFRASTRUCTURE.  Run in the build container only (the GPU box has no /root/reference):

    python -m oracle.make_golden

Every case is fully described by (config, seeds); weights, inputs and noise are regenerated from
the seeds with ``oracle.dreamer_oracle.make_weights / make_noise`` so only the reference's OUTPUTS
are stored (loss scalars, rollout tensors and parameter gradients of a reduced-width model).
"""
from __future__ import annotations

import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import dreamer_oracle as O  # noqa: E402
from oracle import ref_harness as RH  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

# reduced widths keep the stored gradients small; sizes are multiples of 4 as the kernels require
SMALL = dict(S=8, D=32, Hd=32, act_hidden=32, reward_hidden=48, value_hidden=32, pcont_hidden=32)

CASES = {
    # name: (A, dist, use_pcont, T, B, H, sizes)
    "tanh_small": dict(A=3, dist="tanh_normal", use_pcont=False, T=5, B=3, H=4, **SMALL),
    "onehot_pcont_small": dict(A=4, dist="one_hot", use_pcont=True, T=5, B=3, H=4, **SMALL),
    "relaxed_small": dict(A=4, dist="relaxed_one_hot", use_pcont=False, T=4, B=2, H=3, **SMALL),
    # default widths (stoch 30 / deter 200 / hidden 200), cartpole-sized action: losses + rollouts only
    "tanh_default": dict(A=1, dist="tanh_normal", use_pcont=False, T=6, B=4, H=5, S=30, D=200, Hd=200,
                         act_hidden=200, reward_hidden=300, value_hidden=200, pcont_hidden=200),
}


# Full-size cases (BASELINE.json configs[1] and configs[2]: B=50, T=50, H=15, default widths).  The reference runs in
# fp64 (the arbiter of SURVEY 8(c): reference fp32-vs-fp64 is <= 6e-7 on outputs, but fp32 summation noise reaches 2.6e-4
# of |g|max on the smallest world-model gradients); tensors are stored as SUMMARIES (abs-max, sum, a strided sample of
# <= 4096 elements), see `summarize`.
FULL_CASES = {
    "walker_full": dict(A=6, dist="tanh_normal", use_pcont=False, T=50, B=50, H=15, S=30, D=200, Hd=200, act_hidden=200,
                        reward_hidden=300, value_hidden=200, pcont_hidden=200, kl_scale=1.0),
    "pong_full": dict(A=6, dist="one_hot", use_pcont=True, T=50, B=50, H=15, S=30, D=200, Hd=200, act_hidden=200,
                      reward_hidden=300, value_hidden=200, pcont_hidden=200, kl_scale=0.1),
}
SAMPLE = 4096


def sample_stride(numel: int) -> int:
    return max(1, numel // SAMPLE)


def summarize(t: torch.Tensor) -> dict:
    """abs-max, sum and a strided sample of a tensor: enough to evaluate the normwise error max|x - ref| / max|ref| on
    the sample without storing 30 MB tensors."""
    flat = t.detach().reshape(-1).double()
    st = sample_stride(flat.numel())
    return dict(absmax=float(flat.abs().max()), sum=float(flat.sum()), numel=flat.numel(), stride=st,
                sample=flat[::st][:SAMPLE].float().clone())


def summary_err(x: torch.Tensor, ref: dict) -> float:
    """Normwise error of `x` against a stored summary (sample + mean), both scaled by the reference abs-max."""
    flat = x.detach().reshape(-1).double().cpu()
    assert flat.numel() == ref["numel"], (flat.numel(), ref["numel"])
    scale = max(ref["absmax"], 1e-30)
    e_sample = float((flat[::ref["stride"]][:SAMPLE] - ref["sample"].double()).abs().max()) / scale
    e_mean = abs(float(flat.sum()) - ref["sum"]) / (scale * ref["numel"])
    return max(e_sample, e_mean)


def case_inputs(c, dtype=torch.float32):
    """Deterministic weights / inputs / noise of a case (shared with the tests)."""
    shapes = O.weight_shapes(c["A"], c["S"], c["D"], c["Hd"], 1024, c["dist"], c["act_hidden"], 3,
                             c["reward_hidden"], 3, c["value_hidden"], 3, c["use_pcont"], c["pcont_hidden"], 3,
                             conv=True)
    w = O.make_weights(shapes, seed=11, dtype=dtype)
    g = torch.Generator().manual_seed(12)
    T, B, A = c["T"], c["B"], c["A"]
    obs = torch.randint(0, 256, (T + 1, B, 3, 64, 64), generator=g, dtype=torch.uint8)
    if c["dist"] == "tanh_normal":
        act = (torch.rand(T + 1, B, A, generator=g, dtype=torch.float64) * 2 - 1).to(dtype)
    else:
        idx = torch.randint(0, A, (T + 1, B), generator=g)
        act = torch.nn.functional.one_hot(idx, A).to(dtype)
    rew = torch.randn(T + 1, B, generator=g, dtype=torch.float64).to(dtype)
    done = torch.rand(T, B, generator=g) < 0.2
    N = (T - 1) * B if c["use_pcont"] else T * B
    noise = O.make_noise(T, B, N, c["H"], c["S"], A, seed=13, dist=c["dist"], dtype=dtype)
    return w, obs, act, rew, done, noise


def model_kwargs(c):
    return dict(stochastic_size=c["S"], deterministic_size=c["D"], hidden_size=c["Hd"],
                action_hidden_size=c["act_hidden"], reward_hidden=c["reward_hidden"], value_hidden=c["value_hidden"],
                pcont_hidden=c["pcont_hidden"])


def run_reference(c):
    w, obs, act, rew, done, noise = case_inputs(c)
    model = RH.build_reference_model(w, c["A"], c["dist"], c["use_pcont"], **model_kwargs(c))
    # conv weights come from `w` too (build_reference_model copies every key it finds)
    algo, model_loss, actor_loss, value_loss, info = RH.reference_loss(
        model, obs, act, rew, done, noise, c["H"], c["dist"], use_pcont=c["use_pcont"])
    model.zero_grad()
    model_loss.backward()
    actor_loss.backward()
    value_loss.backward()
    out = dict(case=c, model_loss=model_loss.detach(), actor_loss=actor_loss.detach(),
               value_loss=value_loss.detach(), div=info.divergence.detach(), reward_loss=info.reward_loss.detach(),
               image_loss=info.image_loss.detach(), pcont_loss=info.pcont_loss.detach(),
               prior_ent=info.prior_entropy.detach(), post_ent=info.post_entropy.detach())
    store_grads = c["D"] <= 64
    grads = {}
    for k, p in model.named_parameters():
        if p.grad is None:
            continue
        if k.startswith("observation_"):
            continue
        canon = k
        for alias in ("representation._transition_model.", "rollout.representation_model._transition_model.",
                      "rollout.transition_model."):
            if k.startswith(alias):
                canon = "transition." + k[len(alias):]
        if canon.startswith("rollout.representation_model._stochastic_posterior_model."):
            canon = "representation._stochastic_posterior_model." + canon.split("_stochastic_posterior_model.")[1]
        if store_grads:
            grads[canon] = p.grad.detach().clone()
        else:
            grads[canon] = torch.stack([p.grad.abs().max(), p.grad.double().sum().float()])
    out["grads"] = grads
    out["grads_full"] = store_grads
    # rollouts on the conv embedding of the same observations (separate pass: the loss consumed the noise)
    with torch.no_grad():
        embed = model.observation_encoder(obs[:-1].float() / 255.0 - 0.5)
    prior, post, imag, actions = RH.reference_rollouts(model, embed, act[1:], noise, c["H"], c["dist"], c["use_pcont"])
    for name, st in (("prior", prior), ("post", post), ("imag", imag)):
        for f in ("mean", "std", "stoch", "deter"):
            out[f"{name}_{f}"] = getattr(st, f).detach()
    out["actions"] = actions.detach()
    return out


def canonical(k: str) -> str:
    for alias in ("representation._transition_model.", "rollout.representation_model._transition_model.",
                  "rollout.transition_model."):
        if k.startswith(alias):
            return "transition." + k[len(alias):]
    if k.startswith("rollout.representation_model._stochastic_posterior_model."):
        return "representation._stochastic_posterior_model." + k.split("_stochastic_posterior_model.")[1]
    return k


def run_reference_full(c):
    """Full-size case: the reference in fp64, outputs stored as summaries."""
    dt = torch.float64
    # the fp32 inputs every test regenerates, widened exactly
    w, obs, act, rew, done, noise = case_inputs(c)
    w = {k: v.to(dt) for k, v in w.items()}
    act, rew = act.to(dt), rew.to(dt)
    noise = {k: v.to(dt) for k, v in noise.items()}
    model = RH.build_reference_model(w, c["A"], c["dist"], c["use_pcont"], dtype=dt, **model_kwargs(c))
    algo, model_loss, actor_loss, value_loss, info = RH.reference_loss(
        model, obs, act, rew, done, noise, c["H"], c["dist"], use_pcont=c["use_pcont"], kl_scale=c["kl_scale"])
    model.zero_grad()
    model_loss.backward()
    actor_loss.backward()
    value_loss.backward()
    out = dict(case=c, summary=True)
    for k, v in (("model_loss", model_loss), ("actor_loss", actor_loss), ("value_loss", value_loss),
                 ("div", info.divergence), ("reward_loss", info.reward_loss), ("image_loss", info.image_loss),
                 ("pcont_loss", info.pcont_loss), ("prior_ent", info.prior_entropy), ("post_ent", info.post_entropy)):
        out[k] = v.detach().float()
    grads = {}
    for k, p in model.named_parameters():
        if p.grad is not None:
            grads[canonical(k)] = summarize(p.grad)
    out["grads"] = grads
    with torch.no_grad():
        embed = model.observation_encoder(obs[:-1].to(dt) / 255.0 - 0.5)
    prior, post, imag, actions = RH.reference_rollouts(model, embed, act[1:], noise, c["H"], c["dist"], c["use_pcont"])
    for name, st in (("prior", prior), ("post", post), ("imag", imag)):
        for f in ("mean", "std", "stoch", "deter"):
            out[f"{name}_{f}"] = summarize(getattr(st, f))
    out["actions"] = summarize(actions)
    out["embed"] = summarize(embed)
    # The reference's own fp32 run against its fp64 run: the rounding noise of the reference arithmetic itself, per
    # gradient.  With a 12 288-pixel image loss summed over 2 500 frames behind a 50-step recurrence some world-model
    # gradients carry 2e-4 .. 1e-3 of |g|max of fp32 noise -- no fp32 implementation can be closer to the truth than that,
    # so the parity tests use max(1e-4, 4 x this noise) for those tensors.
    w32, obs, act32, rew32, done, noise32 = case_inputs(c)
    m32 = RH.build_reference_model(w32, c["A"], c["dist"], c["use_pcont"], **model_kwargs(c))
    _, ml, al, vl, _ = RH.reference_loss(m32, obs, act32, rew32, done, noise32, c["H"], c["dist"],
                                         use_pcont=c["use_pcont"], kl_scale=c["kl_scale"])
    m32.zero_grad()
    ml.backward(); al.backward(); vl.backward()
    out["fp32_noise"] = {canonical(k): summary_err(p.grad, grads[canonical(k)])
                         for k, p in m32.named_parameters() if p.grad is not None}
    return out


def main():
    assert RH.reference_available(), "needs /root/reference"
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    torch.set_num_threads(4)
    if "--full" in sys.argv:
        torch.set_num_threads(8)
        for name, c in FULL_CASES.items():
            out = run_reference_full(c)
            path = os.path.join(GOLDEN_DIR, f"{name}.pt")
            torch.save(out, path)
            print(f"{name}: model {out['model_loss']:.6f} actor {out['actor_loss']:.6f} value {out['value_loss']:.6f} "
                  f"-> {path} ({os.path.getsize(path) / 1024:.0f} KiB)")
        return
    for name, c in CASES.items():
        out = run_reference(c)
        path = os.path.join(GOLDEN_DIR, f"{name}.pt")
        torch.save(out, path)
        print(f"{name}: model {out['model_loss']:.6f} actor {out['actor_loss']:.6f} value {out['value_loss']:.6f} "
              f"-> {path} ({os.path.getsize(path) / 1024:.0f} KiB)")


if __name__ == "__main__":
    main()
