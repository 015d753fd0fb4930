"""Policy-value network wrapper with the reference's surface (neural_network.py:94-360 there).

`AlphaZeroGomokuNet` is a plain torch module whose state-dict keys are the reference's (45 entries:
conv, bn, residual_tower.{0,1}.{conv1,conv2,bn1,bn2}, policy_conv, policy_fc, value_conv, value_fc1,
value_fc2), so checkpoints interchange and `training.py`-style code can keep calling
`model.model(x)` / `model.model.parameters()`.  Inference on the hot path does not go through
torch: `GomokuModel.predict` hands the position to the CUDA engine (gz_pv_forward), which re-packs
the weights (BatchNorm folded) whenever the parameters change.
"""
import glob
import logging
import os

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

import engine as _engine


class ResidualBlock(nn.Module):
    def __init__(self, channels):
        super().__init__()
        self.conv1 = nn.Conv2d(channels, channels, 3, padding=1)
        self.conv2 = nn.Conv2d(channels, channels, 3, padding=1)
        self.bn1 = nn.BatchNorm2d(channels)
        self.bn2 = nn.BatchNorm2d(channels)

    def forward(self, x):
        y = F.relu(self.bn1(self.conv1(x)))
        return F.relu(self.bn2(self.conv2(y)) + x)


class AlphaZeroGomokuNet(nn.Module):
    """conv3x3(3->128)+BN+ReLU, residual blocks, policy head 128->2->450->225, value head
    128->1->225->64->1 tanh.  Heads have no BN and no ReLU after their 1x1 convs."""

    def __init__(self, board_size=15, channels=3, num_residual=2):
        super().__init__()
        self.board_size, self.channels, self.num_residual = board_size, channels, num_residual
        cells = board_size * board_size
        self.conv = nn.Conv2d(channels, 128, 3, padding=1)
        self.bn = nn.BatchNorm2d(128)
        self.residual_tower = nn.ModuleList(ResidualBlock(128) for _ in range(num_residual))
        self.policy_conv = nn.Conv2d(128, 2, 1)
        self.policy_fc = nn.Linear(2 * cells, cells)
        self.value_conv = nn.Conv2d(128, 1, 1)
        self.value_fc1 = nn.Linear(cells, 64)
        self.value_fc2 = nn.Linear(64, 1)

    def forward(self, x):
        h = F.relu(self.bn(self.conv(x)))
        for block in self.residual_tower:
            h = block(h)
        policy = self.policy_fc(self.policy_conv(h).flatten(1))
        v = F.relu(self.value_fc1(self.value_conv(h).flatten(1)))
        return policy, torch.tanh(self.value_fc2(v))


def weights_version(module):
    """Cheap change detector: torch bumps `_version` on every in-place update (optimizer steps,
    load_state_dict), so the engine re-packs weights only when this tuple changes.  Updates torch cannot see
    (parameters written by a CUDA-graph replay) are announced with mark_weights_changed()."""
    return (id(module), getattr(module, "_gz_generation", 0)) + \
        tuple(int(t._version) for t in module.state_dict(keep_vars=True).values())


def mark_weights_changed(module):
    """Call after updating a module's parameters behind torch's back (training.GraphedTrainStep replays)."""
    module._gz_generation = getattr(module, "_gz_generation", 0) + 1


def planes_to_cells(board_state):
    """(15,15) ints or (3,15,15) planes -> uint8[225] cells (neural_network.py:224-231 accepts both)."""
    a = np.asarray(board_state)
    if a.ndim == 2:
        return np.ascontiguousarray(a, dtype=np.uint8).reshape(225)
    return (np.asarray(a[0] > 0.5, np.uint8) + 2 * np.asarray(a[1] > 0.5, np.uint8)).reshape(225)


class GomokuModel:
    """Model wrapper: create, predict (through the CUDA engine), save, load, auto-discover."""

    BEST = os.path.join("models", "alphazero_gomoku_best.pth")

    def __init__(self, model_path=None, board_size=15, device="cuda"):
        self.board_size = board_size
        self.model_path = model_path
        # "cpu"/None requests are served by the CUDA engine (no CPU inference path is built); without
        # a GPU the torch module still constructs (checkpoints, state dicts) but predict() raises.
        want = "cuda" if device in (None, "cpu") else device
        self.device = torch.device(want if torch.cuda.is_available() else "cpu")
        self.logger = logging.getLogger(__name__)
        self.model = AlphaZeroGomokuNet(board_size).to(self.device)
        self.model.eval()
        self._engine = None
        self._packed = None
        self._logged_no_pretrained_once = False
        # load order of the reference: explicit path, ./models best, newest ./models/alphazero_gomoku_*
        if model_path and os.path.exists(model_path):
            self.load_model(model_path)
        elif os.path.exists(self.BEST):
            self.load_model(self.BEST)
        else:
            latest = self._find_latest_model()
            if latest:
                self.load_model(latest)
            elif not self._logged_no_pretrained_once:
                self.logger.info("No pretrained model found, using random initialization")
                self._logged_no_pretrained_once = True

    # ---- engine plumbing
    def _get_engine(self):
        if self._engine is None:
            self._engine = _engine.Engine(max_games=64, max_simulations=8,
                                          device=self.device.index or 0)
        v = weights_version(self.model)
        if v != self._packed:
            self._engine.set_pv_weights(self.model.state_dict())
            self._packed = v
        return self._engine

    def predict_batch(self, cells, to_move=None, move_count=None):
        """cells uint8[G,225] -> (policy float32[G,225], value float32[G]) on the host."""
        cells = np.ascontiguousarray(cells, np.uint8).reshape(-1, 225)
        eng = self._get_engine()
        out_p, out_v = [], []
        for i in range(0, len(cells), eng.max_games):
            c = cells[i:i + eng.max_games]
            stones = (c != 0).sum(1)
            boards = eng.boards_from_cells(c, 1 + stones % 2, stones)
            _, probs, value = eng.pv_forward(boards)
            out_p.append(probs.cpu().numpy())
            out_v.append(value.cpu().numpy())
        return np.concatenate(out_p), np.concatenate(out_v)

    def predict(self, board_state):
        """(policy float32[225] softmax over all cells, value float) for one position."""
        policy, value = self.predict_batch(planes_to_cells(board_state)[None])
        return policy[0], float(value[0])

    def get_move_probabilities(self, board_state, valid_moves):
        policy, _ = self.predict(board_state)
        probs = np.array([policy[r * self.board_size + c] for r, c in valid_moves], dtype=np.float64)
        total = probs.sum()
        return probs / total if total > 0 else probs

    # ---- checkpoints
    def _find_latest_model(self):
        files = glob.glob(os.path.join("models", "alphazero_gomoku_*.pth"))
        return max(files, key=os.path.getmtime) if files else None

    def save_model(self, filepath):
        folder = os.path.dirname(filepath)
        if folder:
            os.makedirs(folder, exist_ok=True)
        torch.save({"model_state_dict": self.model.state_dict(), "model_type": "alphazero_gomoku",
                    "board_size": self.board_size, "device": str(self.device)}, filepath)
        self.logger.info("Model saved to %s", filepath)

    def load_model(self, filepath):
        if not os.path.exists(filepath):
            self.logger.warning("Model file not found: %s", filepath)
            return False
        try:
            ckpt = torch.load(filepath, map_location=self.device)
            missing, unexpected = self.model.load_state_dict(ckpt["model_state_dict"], strict=False)
            self.model.eval()
            self._packed = None
            if missing or unexpected:
                self.logger.warning("Loaded with mismatches. missing=%d, unexpected=%d",
                                    len(missing), len(unexpected))
            return True
        except Exception as exc:           # the reference downgrades every load error to a warning
            self.logger.warning("Failed to load model: %s", exc)
            return False

    def auto_save_model(self, iteration=None, suffix=""):
        name = ("alphazero_gomoku_iter_%s%s.pth" % (iteration, suffix) if iteration is not None
                else "alphazero_gomoku_auto%s.pth" % suffix)
        path = os.path.join("models", name)
        self.save_model(path)
        return path

    def train_mode(self):
        self.model.train()

    def eval_mode(self):
        self.model.eval()
