"""Import shim: put reth_b200/compat on sys.path (ahead of the reference's reth_buffer) and
`import reth_buffer` resolves to the B200 engine with the reference's exported names
(reth_buffer/reth_buffer/__init__.py:5-8), so test/apex-dqn/{trainer,worker}.py and
reth_buffer/test/test_rb.py run unmodified."""
from reth_b200.sampler import PERSampler  # noqa: F401
from reth_b200.service import Client, NumpyLoader, TorchCudaLoader, start_per, start_server  # noqa: F401
