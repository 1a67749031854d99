"""agilerl_b200 — B200-native (sm_100a) experience-path engine with AgileRL's components API.

``from agilerl_b200.components import ReplayBuffer, MultiStepReplayBuffer,
PrioritizedReplayBuffer, RolloutBuffer, Sampler`` mirrors ``agilerl.components``;
``agilerl_b200.install()`` swaps the classes into an importable ``agilerl`` package.
The compute path is libb200xp.so (include/b200xp.h); there is no CPU fallback.
"""

__version__ = "0.1.0"


def install() -> None:
    """Monkey-patch ``agilerl.components`` so existing AgileRL code uses the B200 engine."""
    import importlib

    from . import components as c

    mods = {
        "agilerl.components": ("ReplayBuffer", "MultiStepReplayBuffer", "PrioritizedReplayBuffer",
                               "MultiAgentReplayBuffer", "RolloutBuffer", "Sampler"),
        "agilerl.components.multi_agent_replay_buffer": ("MultiAgentReplayBuffer",),
        "agilerl.components.replay_buffer": ("ReplayBuffer", "MultiStepReplayBuffer", "PrioritizedReplayBuffer"),
        "agilerl.components.segment_tree": ("SumSegmentTree", "MinSegmentTree"),
        "agilerl.components.rollout_buffer": ("RolloutBuffer",),
        "agilerl.components.sampler": ("Sampler",),
    }
    for name, symbols in mods.items():
        mod = importlib.import_module(name)
        for s in symbols:
            setattr(mod, s, getattr(c, s))


def link_n_step(memory, n_step_memory) -> None:
    """Optional one-liner for the off-policy loop (training/train_off_policy.py:46-60 creates ``memory`` and
    ``n_step_memory`` side by side): ``agilerl_b200.link_n_step(memory, n_step_memory)`` lets
    ``memory.sample(batch_size, beta)`` gather the rows of both buffers in one launch, so that the
    reference's next call ``n_step_sampler.sample(experiences["idxs"])`` costs no second launch.  Same
    results with or without it (``PrioritizedReplayBuffer.link_n_step``)."""
    memory.link_n_step(n_step_memory)
