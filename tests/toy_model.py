"""CPU stand-in for the model with its calling convention (host-logic tests of the engine; never on the product path)."""
import torch


class ToyModel:
    """next token = (last input + 1) mod V, elapsed += T; records every call and leaves a trace of what each slot consumed."""

    def __init__(self, V=50, L=2, C=8):
        self.device = torch.device("cpu")
        self.V, self.L, self.C = V, L, C
        self.vocab_size = V
        self.calls = []

    def generate_zero_state(self, bsz):
        return [torch.zeros((self.L, 2, bsz, self.C), dtype=torch.float16), torch.zeros((self.L, bsz, 1, 64, 64), dtype=torch.float16),
                torch.zeros((bsz,), dtype=torch.int32)]

    def _tokens(self, idxs):
        return torch.tensor(idxs, dtype=torch.int32).reshape(-1)

    def _step(self, tokens, B, T, state, full_output, want_logits=True, slot_idx=None, out=None):
        self.calls.append((B, T, slot_idx.tolist() if slot_idx is not None else None))
        tk = tokens.view(B, T).long()
        state[2] += T
        rows = slot_idx.long()
        state[0][0, 0, rows, 0] += tk.sum(1).to(torch.float16)
        if not want_logits:
            return None
        lg = torch.zeros((B, self.V), dtype=torch.float16)
        lg[torch.arange(B), (tk[:, -1] + 1) % self.V] = 5.0
        return lg


def make_toy_model():
    return ToyModel()
