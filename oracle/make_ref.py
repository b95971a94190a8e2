#!/usr/bin/env python
"""Build-time recipe for oracle/_ref/: the reference's OWN Python modules for the replay hot path, taken from
/root/reference where they lie and placed (unchanged, byte for byte) under oracle/_ref/reth_ref/ so that they
travel to the GPU box with the snapshot -- the reference checkout itself does not exist there.  oracle/_ref/ is
git-ignored: no reference source enters the history.  Used ONLY by bench.py's CPU legs (cpu_baseline /
--impl reference) as the CPU baseline of record, SURVEY.md 8d "Tier A":

    FIFOPolicy.get_indices   reth_buffer/reth_buffer/cache_policy/fifo_policy.py:11-18
    NumpyBuffer.append_batch reth/reth/buffer/buffer.py:57-86     .select :92-93
    PERSampler.update/sample reth_buffer/reth_buffer/sampler/per_sampler.py:24-35
    NumbaSumTree (njit)      reth_buffer/reth_buffer/utils/sumtree.py:5-113
    Schedule                 reth_buffer/reth_buffer/utils/schedule.py:4-52

The package __init__ files written here are this repo's own two-liners: the reference's import its whole service
(zmq / lmdb / lz4 / portpicker), which this path does not need and this image does not have.
Run by __graft_entry__.build() when /root/reference is present."""
import os
import shutil

REF = os.environ.get("RETH_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref", "reth_ref")
FILES = {
    "reth_buffer/reth_buffer/utils/sumtree.py": "utils/sumtree.py",
    "reth_buffer/reth_buffer/utils/schedule.py": "utils/schedule.py",
    "reth_buffer/reth_buffer/sampler/per_sampler.py": "sampler/per_sampler.py",
    "reth_buffer/reth_buffer/sampler/base_sampler.py": "sampler/base_sampler.py",
    "reth_buffer/reth_buffer/cache_policy/fifo_policy.py": "cache_policy/fifo_policy.py",
    "reth_buffer/reth_buffer/cache_policy/base_policy.py": "cache_policy/base_policy.py",
    "reth/reth/buffer/buffer.py": "buffer/buffer.py",
}
INITS = {
    "": "",
    "utils": "from .sumtree import NumbaSumTree  # noqa: F401\nfrom .schedule import Schedule  # noqa: F401\n",
    "sampler": "from .per_sampler import PERSampler  # noqa: F401\n",
    "cache_policy": "from .fifo_policy import FIFOPolicy  # noqa: F401\n",
    "buffer": "from .buffer import NumpyBuffer  # noqa: F401\n",
}


def available():
    return all(os.path.exists(os.path.join(REF, p)) for p in FILES)


def build(force=False):
    """-> path of oracle/_ref (to put on sys.path) or None when the reference checkout is absent."""
    if not available():
        return os.path.dirname(DST) if os.path.exists(os.path.join(DST, "utils", "sumtree.py")) else None
    for src, dst in FILES.items():
        out = os.path.join(DST, dst)
        os.makedirs(os.path.dirname(out), exist_ok=True)
        if force or not os.path.exists(out) or os.path.getmtime(out) < os.path.getmtime(os.path.join(REF, src)):
            shutil.copyfile(os.path.join(REF, src), out)
    for pkg, body in INITS.items():
        with open(os.path.join(DST, pkg, "__init__.py"), "w") as f:
            f.write(body)
    return os.path.dirname(DST)


if __name__ == "__main__":
    print(build(force=True))
