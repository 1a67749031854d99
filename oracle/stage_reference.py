"""TEST / BASELINE INFRASTRUCTURE — stages the UNMODIFIED reference sources the oracle loader needs.

``/root/reference`` exists only in the build container.  ``bench.py --impl reference`` and the
``cpu_baseline`` legs must time the reference's own Python on the GPU box's host cores, so ``build()``
(``__graft_entry__.py``) calls ``stage()`` here: the few files ``oracle/ref_loader.py`` loads by path are
copied byte for byte into ``oracle/_ref/agilerl/...`` (git-ignored: never in history; not gpurun-ignored:
it travels to the GPU box like the built ``.so`` files) together with ``MANIFEST.json`` (sha256 per file).
``ref_loader`` reads ``/root/reference`` when it exists and ``oracle/_ref`` otherwise, and refuses a staged
file whose hash does not match the manifest.  Nothing is edited; nothing in ``agilerl_b200/`` reads this.
"""

from __future__ import annotations

import hashlib
import json
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
STAGE_ROOT = os.path.join(HERE, "_ref")
SOURCE_ROOT = os.environ.get("AGILERL_REFERENCE", "/root/reference")

# every path oracle/ref_loader.py opens (relative to the reference root)
FILES = [
    "agilerl/components/segment_tree.py",
    "agilerl/components/replay_buffer.py",
    "agilerl/components/rollout_buffer.py",
    "agilerl/components/multi_agent_replay_buffer.py",
    "agilerl/typing.py",
    "agilerl/utils/algo_utils.py",
    "agilerl/utils/llm_utils.py",
    "agilerl/algorithms/ppo.py",
    "agilerl/algorithms/ippo.py",
    "agilerl/algorithms/ppo_llm.py",
]


def _sha(path: str) -> str:
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for chunk in iter(lambda: f.read(1 << 20), b""):
            h.update(chunk)
    return h.hexdigest()


def stage(verbose: bool = False) -> str | None:
    """Copy FILES from the reference tree into oracle/_ref (no-op, returning None, where the reference
    tree is absent — e.g. on the GPU box, which uses what was staged in the build container)."""
    if not os.path.isfile(os.path.join(SOURCE_ROOT, FILES[0])):
        return None
    manifest = {}
    for rel in FILES:
        src, dst = os.path.join(SOURCE_ROOT, rel), os.path.join(STAGE_ROOT, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)
        manifest[rel] = _sha(dst)
        assert manifest[rel] == _sha(src)
    with open(os.path.join(STAGE_ROOT, "MANIFEST.json"), "w") as f:
        json.dump({"source": SOURCE_ROOT, "sha256": manifest}, f, indent=1, sort_keys=True)
    if verbose:
        print(f"staged {len(FILES)} unmodified reference files under {STAGE_ROOT}")
    return STAGE_ROOT


def verify(root: str = STAGE_ROOT) -> bool:
    """True when every staged file matches the manifest written at staging time."""
    try:
        with open(os.path.join(root, "MANIFEST.json")) as f:
            manifest = json.load(f)["sha256"]
    except OSError:
        return False
    return all(os.path.isfile(os.path.join(root, rel)) and _sha(os.path.join(root, rel)) == sha
               for rel, sha in manifest.items()) and set(manifest) == set(FILES)


if __name__ == "__main__":
    print(stage(verbose=True), verify())
