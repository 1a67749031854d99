"""Sampler with the dispatch rules of agilerl/components/sampler.py (:53-113): ``.sample`` is
bound to standard / prioritized / n-step / distributed sampling by ``isinstance`` of the memory."""

from __future__ import annotations

import warnings
from typing import Any

import torch
from torch.utils.data import DataLoader

from .data import ReplayDataset
from .replay_buffer import MultiStepReplayBuffer, PrioritizedReplayBuffer, ReplayBuffer
from .tensor_batch import TensorBatch


class Sampler:
    @staticmethod
    def tensordict_collate_fn(batch: list) -> TensorBatch:
        return TensorBatch({key: torch.stack([b[key] for b in batch]) for key in batch[0].keys()},
                           batch_size=len(batch))

    def __init__(self, memory=None, dataset: ReplayDataset | None = None,
                 dataloader: DataLoader | None = None) -> None:
        assert (memory is not None) or ((dataset is not None) and (dataloader is not None)), (
            "Sampler needs to be initialized with either 'memory' or ('dataset' AND 'dataloader').")
        self.distributed = (dataloader is not None and dataset is not None
                            and isinstance(dataloader, DataLoader))
        self.per = isinstance(memory, PrioritizedReplayBuffer)
        self.n_step = isinstance(memory, MultiStepReplayBuffer)
        self.memory = memory
        self.dataset = dataset
        if self.distributed:
            if isinstance(dataset.buffer, ReplayBuffer):
                self.dataloader = self._replace_dataloader_collate_fn(dataloader)
            else:
                self.dataloader = dataloader
            if not isinstance(self.dataset, ReplayDataset):
                warnings.warn("Dataset is not an agilerl ReplayDataset.", stacklevel=2)
            self.sample = self.sample_distributed
        else:
            self.dataloader = dataloader
            if self.per:
                self.sample = self.sample_per
            elif self.n_step:
                self.sample = self.sample_n_step
            else:
                if not isinstance(self.memory, ReplayBuffer) and not hasattr(self.memory, "sample"):
                    warnings.warn("Memory is not an agilerl ReplayBuffer or MultiAgentReplayBuffer.",
                                  stacklevel=2)
                self.sample = self.sample_standard

    def _replace_dataloader_collate_fn(self, dataloader: DataLoader) -> DataLoader:
        params = {"dataset": dataloader.dataset, "batch_size": dataloader.batch_size,
                  "collate_fn": self.tensordict_collate_fn}
        if dataloader.num_workers != 0:
            params["num_workers"] = dataloader.num_workers
        if dataloader.pin_memory:
            params["pin_memory"] = True
        if dataloader.drop_last:
            params["drop_last"] = True
        return DataLoader(**params)

    def sample_standard(self, batch_size: int, return_idx: bool = False):
        return self.memory.sample(batch_size, return_idx)

    def sample_distributed(self, batch_size: int, return_idx: bool | None = None):
        self.dataset.batch_size = batch_size
        return next(iter(self.dataloader))

    def sample_per(self, batch_size: int, beta: float):
        return self.memory.sample(batch_size, beta)

    def sample_n_step(self, idxs: Any):
        return self.memory.sample_from_indices(idxs)

    @classmethod
    def create_dataloader(cls, dataset: ReplayDataset, batch_size: int | None = None, **kwargs) -> DataLoader:
        if "collate_fn" not in kwargs:
            kwargs["collate_fn"] = cls.tensordict_collate_fn
        return DataLoader(dataset, batch_size=batch_size, **kwargs)
