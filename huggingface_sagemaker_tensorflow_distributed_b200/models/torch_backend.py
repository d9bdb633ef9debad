"""PyTorch-autograd execution backend: the CPU path (BASELINE config #1 runs
``singe_node_train.py`` on CPU) and the fp32 oracle the B200 engine is tested against.
Gradient averaging here is a library all-reduce (gloo on CPU); the hand-written NVLink
kernels live in the engine backend."""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch
import torch.nn.functional as F

from ..data import IGNORE_LABEL
from ..optim import Adam, KerasAdamW
from ..parallel import comm
from .configuration import EncoderConfig
from .reference import EncoderForSequenceClassification


class TorchBackend:
    name = "torch"

    def __init__(self, cfg: EncoderConfig, device: torch.device, dtype: str = "fp32", seed: int = 42):
        self.cfg, self.device = cfg, device
        self.autocast = dtype == "bf16"
        gen_state = torch.random.get_rng_state()
        torch.manual_seed(seed)
        self.model = EncoderForSequenceClassification(cfg).to(device)
        torch.random.set_rng_state(gen_state)
        self.optimizer: Optional[KerasAdamW] = None
        self.distributed = False
        self._pending_opt_state: Optional[dict] = None

    # ---- state ------------------------------------------------------------------
    def state_dict(self) -> Dict[str, torch.Tensor]:
        return {k: v.detach().float().cpu() for k, v in self.model.state_dict().items()}

    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        own = self.model.state_dict()
        for k, v in sd.items():
            if k in own:
                own[k].copy_(v.to(own[k].dtype))

    def optimizer_state(self) -> dict:
        if self.optimizer is None:
            return {}
        names = [n for n, _ in self.model.named_parameters()]
        st = {}
        for n, p in zip(names, self.model.parameters()):
            s = self.optimizer.state.get(p)
            if s:
                st[n] = {"m": s["m"].cpu(), "v": s["v"].cpu()}
        return {"iterations": self.optimizer.iterations, "state": st}

    def load_optimizer_state(self, st: dict) -> None:
        if not st:
            return
        if self.optimizer is None:  # applied as soon as the optimizer is bound (setup_training)
            self._pending_opt_state = st
            return
        self.optimizer.iterations = int(st.get("iterations", 0))
        for n, p in self.model.named_parameters():
            if n in st.get("state", {}):
                self.optimizer.state[p] = {"m": st["state"][n]["m"].to(p.device), "v": st["state"][n]["v"].to(p.device)}

    def broadcast_state(self, root: int = 0) -> None:
        if comm.size() <= 1:
            return
        for t in self.model.state_dict().values():
            comm.broadcast_(t, root)
        if self.optimizer is not None:
            for p in self.model.parameters():
                s = self.optimizer.state.get(p)
                if s:
                    comm.broadcast_(s["m"], root)
                    comm.broadcast_(s["v"], root)

    # ---- training ---------------------------------------------------------------
    def setup_training(self, opt: Adam, total_steps: int, batch_size: int, seq_len: int) -> None:
        self.optimizer = opt.bind(self.model.parameters(), total_steps)
        self.distributed = bool(opt.distributed) and comm.size() > 1
        if self._pending_opt_state is not None:
            st, self._pending_opt_state = self._pending_opt_state, None
            self.load_optimizer_state(st)

    def _forward(self, feats, labels, train: bool):
        ids = feats["input_ids"].to(self.device, non_blocking=True).long()
        mask = feats["attention_mask"].to(self.device, non_blocking=True)
        lab = labels.to(self.device, non_blocking=True).long()
        self.model.train(train)
        with torch.autocast(self.device.type, dtype=torch.bfloat16, enabled=self.autocast):
            logits = self.model(ids, mask)
        logits = logits.float()
        valid = lab != IGNORE_LABEL
        count = valid.sum()
        loss_sum = F.cross_entropy(logits, lab, ignore_index=IGNORE_LABEL, reduction="sum")
        correct = ((logits.argmax(-1) == lab) & valid).sum()
        return loss_sum, correct, count

    def train_step(self, feats, labels) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        assert self.optimizer is not None, "call compile()/setup_training() first"
        self.optimizer.zero_grad(set_to_none=True)
        loss_sum, correct, count = self._forward(feats, labels, True)
        (loss_sum / count.clamp(min=1)).backward()
        if self.distributed:
            grads = [p.grad for p in self.model.parameters() if p.grad is not None]
            flat = torch.cat([g.reshape(-1) for g in grads])
            comm.allreduce_mean_(flat)
            off = 0
            for g in grads:
                g.copy_(flat[off:off + g.numel()].view_as(g))
                off += g.numel()
        self.optimizer.step()
        return loss_sum.detach(), correct.detach(), count.detach()

    @torch.no_grad()
    def eval_step(self, feats, labels):
        return self._forward(feats, labels, False)

    def finish(self) -> None:
        pass
