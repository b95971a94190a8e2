"""reth_buffer.sampler (reth_buffer/reth_buffer/sampler/__init__.py): the sampler plugin classes."""
from reth_b200.sampler import BaseSampler, FIFOSampler, PERSampler, UniformSampler  # noqa: F401
