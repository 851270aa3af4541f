"""ResNet-50 Collective-mode workload: the public training API a rank uses.

    job = ResNetJob(batch=256)            # reads rank/world from the environment
    for images_u8, labels in loader:      # pinned host tensors, uint8 NHWC / int64
        loss = job.train_step(images_u8, labels)
    print(loss.item())

`train_step` stages the host batch through a double-buffered pinned->device
copy on its own stream (overlapping the previous step's compute), runs the
hand-written uint8->bf16 normalise kernel, then replays the captured step.

Reference parity: this is the program a worker pod of
deploy/examples/resnet.yaml:13-19 runs (there: Paddle's train_fleet.py); the
rank/world come from the same PADDLE_TRAINER_ID / PADDLE_TRAINERS_NUM the
reconciler injects (controllers/paddlejob_helper.go:264,319-322).
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from ..models.resnet import resnet50, resnet_tiny, to_training_dtype
from ..ops import optim as pdj_optim
from ..parallel.collective import CollectiveTrainer
from ..parallel.symm import Communicator


def _loss_fn(model, x, y):
    return pdj_optim.softmax_cross_entropy(model(x), y)


class ResNetJob:
    def __init__(self, batch: int = 256, image: int = 224, num_classes: int = 1000, comm: Optional[Communicator] = None,
                 device: Optional[torch.device] = None, lr: float = 0.1, momentum: float = 0.9,
                 weight_decay: float = 1e-4, use_graph: bool = True, tiny: bool = False, seed: int = 0,
                 bucket_bytes: int = 16 << 20, comm_blocks: int = 24, algo: Optional[int] = None,
                 s2d_stem: bool = True, gemm_1x1: Optional[bool] = None):
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        if gemm_1x1 is None:  # stride-1 1x1 convolutions on the tcgen05 GEMM with fused BN statistics
            gemm_1x1 = os.environ.get("PDJ_RESNET_GEMM_1X1", "1") == "1"
        self.gemm_1x1 = gemm_1x1
        self.batch, self.image = batch, image
        torch.manual_seed(seed)  # identical initial weights on every rank
        self.s2d = s2d_stem and image % 2 == 0
        model = (resnet_tiny if tiny else resnet50)(num_classes=num_classes, fused=True, s2d_stem=self.s2d,
                                                    gemm_1x1=gemm_1x1)
        model = to_training_dtype(model.to(self.device), torch.bfloat16)
        model.train()
        self.model = model
        self.comm = comm
        self.rank = comm.rank if comm else 0
        self.world = comm.nranks if comm else 1
        # static device inputs of the captured step
        # (space-to-depth stem: [N, image/2+3, image/2+3, 16] with a zero border, see ops/optim.py)
        in_shape = pdj_optim.s2d_shape(batch, image, image) if self.s2d else (batch, image, image, 3)
        self._x_nhwc = torch.zeros(in_shape, dtype=torch.bfloat16, device=self.device)
        self.x = self._x_nhwc.permute(0, 3, 1, 2)
        self.y = torch.zeros((batch,), dtype=torch.int64, device=self.device)
        self.trainer = CollectiveTrainer(model, _loss_fn, (self.x, self.y), comm=comm, lr=lr, momentum=momentum,
                                         weight_decay=weight_decay, use_graph=use_graph, bucket_bytes=bucket_bytes,
                                         comm_blocks=comm_blocks, algo=algo, extra_kernels_per_step=1)
        # host -> device staging (double buffered)
        self._stage_u8 = [torch.empty((batch, image, image, 3), dtype=torch.uint8, device=self.device)
                          for _ in range(2)]
        self._stage_y = [torch.empty((batch,), dtype=torch.int64, device=self.device) for _ in range(2)]
        self._copy_stream = torch.cuda.Stream(self.device)
        self._copied = [torch.cuda.Event() for _ in range(2)]
        self._consumed = [torch.cuda.Event() for _ in range(2)]
        for e in self._consumed:
            e.record()
        self._slot = 0
        self._loss_host = torch.zeros(64, dtype=torch.float32).pin_memory()
        self._loss_i = 0
        self.step_count = 0

    @classmethod
    def from_env(cls, arena_bytes: int = 384 << 20, **kw) -> "ResNetJob":
        world = int(os.environ.get("WORLD_SIZE", os.environ.get("PADDLE_TRAINERS_NUM", "1")))
        comm = Communicator.from_env(arena_bytes) if world > 1 else None
        return cls(comm=comm, **kw)

    def prepare(self) -> None:
        """Fills the static inputs with noise and captures the step graph."""
        with torch.no_grad():
            if self.s2d:  # noise in the interior only: the border is the convolution's zero padding
                self._x_nhwc[:, 2:-1, 2:-1, :12].copy_(
                    torch.randn((self.batch, self.image // 2, self.image // 2, 12), device=self.device).to(torch.bfloat16))
            else:
                self.x.copy_(torch.randn(self.x.shape, device=self.device).to(torch.bfloat16))
            self.y.copy_(torch.randint(0, 1000 if self.model.fc.out_features >= 1000 else self.model.fc.out_features,
                                       (self.batch,), device=self.device))
        self.trainer.capture()

    # device-resident step (inputs already in self.x / self.y)
    def step_device(self) -> torch.Tensor:
        self.step_count += 1
        return self.trainer.step()

    def train_step(self, images_u8: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        """One end-to-end step from pinned host tensors.  Returns a pinned host
        scalar view that holds this step's loss once the stream has drained
        (``torch.cuda.current_stream().synchronize()`` or ``.item()`` later)."""
        s = self._slot
        self._slot ^= 1
        main = torch.cuda.current_stream()
        with torch.cuda.stream(self._copy_stream):
            self._copy_stream.wait_event(self._consumed[s])
            self._stage_u8[s].copy_(images_u8, non_blocking=True)
            self._stage_y[s].copy_(labels, non_blocking=True)
            self._copied[s].record(self._copy_stream)
        main.wait_event(self._copied[s])
        if self.s2d:
            pdj_optim.preprocess_u8_s2d(self._stage_u8[s], out=self._x_nhwc)
        else:
            pdj_optim.preprocess_u8_nhwc(self._stage_u8[s], out=self._x_nhwc)
        self.y.copy_(self._stage_y[s])
        self._consumed[s].record(main)
        loss = self.step_device()
        i = self._loss_i
        self._loss_i = (i + 1) % self._loss_host.numel()
        out = self._loss_host[i:i + 1]
        out.copy_(loss.reshape(1), non_blocking=True)
        return out

    @property
    def h2d_bytes_per_step(self) -> int:
        return self.batch * self.image * self.image * 3 + self.batch * 8

    @property
    def d2h_bytes_per_step(self) -> int:
        return 4

    @property
    def launches_per_step(self) -> int:
        return self.trainer.kernels_per_step
