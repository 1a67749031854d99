"""Drop-in for ``agilerl.components`` (agilerl/components/__init__.py:1-13) on the B200 engine."""

from .data import ReplayDataset, Transition
from .multi_agent_replay_buffer import MultiAgentReplayBuffer
from .replay_buffer import MultiStepReplayBuffer, PrioritizedReplayBuffer, ReplayBuffer
from .rollout_buffer import BPTTSequenceType, RolloutArena, RolloutBuffer
from .sampler import Sampler
from .segment_tree import MinSegmentTree, SumSegmentTree
from .tensor_batch import TensorBatch

__all__ = [
    "BPTTSequenceType", "MinSegmentTree", "MultiAgentReplayBuffer", "MultiStepReplayBuffer", "PrioritizedReplayBuffer",
    "ReplayBuffer", "ReplayDataset", "RolloutArena", "RolloutBuffer", "Sampler", "SumSegmentTree",
    "TensorBatch", "Transition",
]
