"""Deterministic synthetic weights shared by the golden generator and the tests.

The reference initialises its three networks from torch's global RNG
(/root/reference/neural_network.py:185, bg_planner.py:210-211).  Fixtures must not depend on
torch's RNG stream, and must not ship 5 MB of floats, so both the generator
(tests/golden/make_goldens.py, which loads these tensors into the *reference* modules) and the
tests (which hand them to the engine) rebuild the same tensors from numpy's PCG64 stream, which is
stable across numpy versions and machines.

Tensor names and shapes are the reference state-dict layout
(neural_network.py:115-130 -> 45 entries, bg_planner.py:34-48 -> 22, bg_planner.py:68-72 -> 6).
BatchNorm running stats and affine terms are randomised on purpose: at PyTorch's default init BN is
almost the identity and would hide folding bugs.
"""
import numpy as np

PV_LAYOUT = [("conv.weight", (128, 3, 3, 3)), ("conv.bias", (128,)), ("bn", 128)]
for _b in range(2):
    _p = "residual_tower.%d." % _b
    PV_LAYOUT += [(_p + "conv1.weight", (128, 128, 3, 3)), (_p + "conv1.bias", (128,)),
                  (_p + "conv2.weight", (128, 128, 3, 3)), (_p + "conv2.bias", (128,)),
                  (_p + "bn1", 128), (_p + "bn2", 128)]
PV_LAYOUT += [("policy_conv.weight", (2, 128, 1, 1)), ("policy_conv.bias", (2,)),
              ("policy_fc.weight", (225, 450)), ("policy_fc.bias", (225,)),
              ("value_conv.weight", (1, 128, 1, 1)), ("value_conv.bias", (1,)),
              ("value_fc1.weight", (64, 225)), ("value_fc1.bias", (64,)),
              ("value_fc2.weight", (1, 64)), ("value_fc2.bias", (1,))]

GRAPHNET_LAYOUT = [("embed.weight", (64, 3, 3, 3)), ("embed.bias", (64,))]
for _i in range(8):
    _k = 3 if _i % 2 == 0 else 1
    GRAPHNET_LAYOUT += [("layers.%d.weight" % _i, (64, 64, _k, _k)), ("layers.%d.bias" % _i, (64,))]
GRAPHNET_LAYOUT += [("policy_head.0.weight", (2, 64, 1, 1)), ("policy_head.0.bias", (2,)),
                    ("policy_head.2.weight", (225, 450)), ("policy_head.2.bias", (225,))]

DQN_LAYOUT = [("net.0.weight", (256, 675)), ("net.0.bias", (256,)),
              ("net.2.weight", (256, 256)), ("net.2.bias", (256,)),
              ("net.4.weight", (225, 256)), ("net.4.bias", (225,))]

LAYOUTS = {"pv": PV_LAYOUT, "graphnet": GRAPHNET_LAYOUT, "dqn": DQN_LAYOUT}


# Default spread per network: the planner nets are scaled up so that GraphNet's softmax and the DQN
# output both move the blended score (bg_planner.py:261) instead of being near-constant.
GAINS = {"pv": 1.0, "graphnet": 2.5, "dqn": 1.5}


def make_state_dict(kind, seed, gain=None):
    """name -> float32 ndarray (int64 for num_batches_tracked), in state-dict order.

    Weights/biases are U(-b, b) with b = gain/sqrt(fan_in) (the scale of PyTorch's default init),
    drawn in float64 from PCG64 and rounded to float32.  `gain` > 1 spreads the logits so that
    argmax/softmax tests are not run on an almost-uniform distribution.
    """
    if gain is None:
        gain = GAINS[kind]
    rng = np.random.default_rng([int(seed), {"pv": 1, "graphnet": 2, "dqn": 3}[kind]])
    out = {}
    last_fan_in = 1
    for name, shape in LAYOUTS[kind]:
        if isinstance(shape, int):                      # a BatchNorm2d group
            c = shape
            out[name + ".weight"] = rng.uniform(0.6, 1.4, c).astype(np.float32)
            out[name + ".bias"] = rng.uniform(-0.3, 0.3, c).astype(np.float32)
            out[name + ".running_mean"] = rng.uniform(-0.4, 0.4, c).astype(np.float32)
            out[name + ".running_var"] = rng.uniform(0.5, 1.8, c).astype(np.float32)
            out[name + ".num_batches_tracked"] = np.array(7, dtype=np.int64)
            continue
        if name.endswith(".weight"):
            last_fan_in = int(np.prod(shape[1:]))
        b = gain / np.sqrt(last_fan_in)
        out[name] = rng.uniform(-b, b, shape).astype(np.float32)
    return out
