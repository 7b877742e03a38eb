"""Shared helpers for the parity tests: seeded synthetic trajectories and error metrics."""
import torch


def rel_err(a, b):
    """L2 relative error ||a - b|| / ||b|| (the tolerance definition of SURVEY.md §8d)."""
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    nb = b.norm().item()
    if nb < 1e-12:
        return (a - b).norm().item()
    return ((a - b).norm() / nb).item()


def make_inputs(B, K, sd, ad, E, seed, pad=True, onehot_actions=False):
    """states N(0,1), actions U(-1,1) (or one-hot), decreasing rtg, timesteps start+arange, LEFT padding on half the batch with the
    upstream pad values (states 0, actions -10, rtg 0, timesteps 0, mask 0)."""
    g = torch.Generator().manual_seed(seed)
    states = torch.randn(B, K, sd, generator=g)
    if onehot_actions:
        actions = torch.nn.functional.one_hot(torch.randint(0, ad, (B, K), generator=g), ad).float()
    else:
        actions = torch.rand(B, K, ad, generator=g) * 2 - 1
    start_r = torch.rand(B, 1, generator=g) * 3.1 + 0.5
    rtg = (start_r - torch.cumsum(torch.rand(B, K, generator=g) * 0.01, dim=1)).unsqueeze(-1)
    start = torch.randint(0, max(E - K, 1), (B, 1), generator=g)
    timesteps = (start + torch.arange(K)[None]).clamp_(max=E - 1)
    mask = torch.ones(B, K)
    if pad:
        for b in range(0, B, 2):
            p = int(torch.randint(0, K // 2 + 1, (1,), generator=g))
            mask[b, :p] = 0
            states[b, :p] = 0
            actions[b, :p] = -10.0
            rtg[b, :p] = 0
            timesteps[b, :p] = 0
    return states, actions, rtg, timesteps, mask


CONFIGS = {
    # name: (state_dim, act_dim, K, hidden, n_layer, n_head, activation, action_tanh)   — SURVEY.md §8 config table
    "c1": dict(state_dim=11, act_dim=3, K=20, hidden_size=128, n_layer=3, n_head=1, activation_function="relu", action_tanh=True),
    "c2": dict(state_dim=1024, act_dim=4, K=30, hidden_size=128, n_layer=3, n_head=1, activation_function="relu", action_tanh=False),
    "c3": dict(state_dim=1024, act_dim=18, K=50, hidden_size=256, n_layer=6, n_head=8, activation_function="relu", action_tanh=False),
    "c4": dict(state_dim=11, act_dim=3, K=200, hidden_size=768, n_layer=12, n_head=12, activation_function="gelu_new", action_tanh=True),
}


def build_pair(name, device, max_ep_len=1000, seed=0, pdrop=0.0):
    """(oracle on CPU, drop-in on `device`) with identical weights."""
    from mtrl_b200.model import DecisionTransformer
    from oracle.decision_transformer import DecisionTransformer as Oracle

    cfg = dict(CONFIGS[name])
    K = cfg.pop("K")
    torch.manual_seed(seed)
    oracle = Oracle(max_length=K, max_ep_len=max_ep_len, resid_pdrop=pdrop, attn_pdrop=pdrop, **cfg)
    # non-trivial LayerNorm / bias values so every gradient path is exercised
    with torch.no_grad():
        for n, p in oracle.named_parameters():
            if n.endswith(".bias"):
                p.normal_(0, 0.02)
            elif "ln_" in n or "embed_ln" in n:
                p.add_(torch.randn_like(p) * 0.05)
    model = DecisionTransformer(max_length=K, max_ep_len=max_ep_len, resid_pdrop=pdrop, attn_pdrop=pdrop, device=device, **cfg)
    model.load_state_dict(oracle.state_dict())
    return oracle, model, K


RELU_MASK_SENSITIVE = ("mlp.c_fc.weight", "mlp.c_fc.bias", "ln_2.weight", "ln_2.bias")


def grad_tolerances(oracle, step_fn, base_tol=2e-2):
    """Per-tensor gradient tolerance.  2e-2 (north star) everywhere, except that with ReLU the tensors directly behind the
    activation mask (mlp.c_fc.{weight,bias}, ln_2.{weight,bias}) inherit sign flips of near-zero pre-activations under ANY bf16 matmul: a
    flipped fraction f of mask entries perturbs a noise-dominated small-batch gradient by ~sqrt(f).  For those the bound is
    max(2e-2, 2 x the error of torch's own bf16 autocast of the oracle on the same inputs) — i.e. "no worse than the library
    bf16 path" (measured: c2 B=8 4.2 %, c2 B=64 2.6 %, c3 B=4 5.4 % for autocast)."""
    import copy

    if oracle.config["activation_function"] != "relu":
        return lambda name: base_tol
    ref = {n: p.grad.clone() for n, p in oracle.named_parameters() if p.grad is not None}
    o2 = copy.deepcopy(oracle)
    o2.zero_grad()
    with torch.autocast("cpu", dtype=torch.bfloat16):
        loss = step_fn(o2)
    loss.float().backward()
    ac = {n: rel_err(p.grad, ref[n]) for n, p in o2.named_parameters() if p.grad is not None and n in ref}

    def tol(name):
        if name.endswith(RELU_MASK_SENSITIVE):
            return max(base_tol, 2.0 * ac.get(name, 0.0))
        return base_tol

    return tol
