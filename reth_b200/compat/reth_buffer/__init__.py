"""Import shim: put reth_b200/compat on sys.path (ahead of the reference's reth_buffer) and
`import reth_buffer` resolves to the B200 engine with the reference's exported names
(reth_buffer/reth_buffer/__init__.py:5-8), so test/apex-dqn/{trainer,worker}.py and
reth_buffer/test/test_rb.py run unmodified.

start_server / start_per spawn the server PROCESS like the reference does (__init__.py:11-27), so
the returned address works from any process (trainer.py:49-66).  RETH_B200_SERVER=inproc keeps the
engine in the calling process instead (learner co-located with the replay: no sockets, the batch
is handed over as a pointer).  Client / NumpyLoader / TorchCudaLoader pick the transport from the
address."""
import os

from reth_b200.sampler import PERSampler  # noqa: F401
from reth_b200.service import Client, NumpyLoader, TorchCudaLoader  # noqa: F401
from reth_b200 import service as _inproc
from reth_b200 import service_mp as _mp


def _mod():
    return _inproc if os.environ.get("RETH_B200_SERVER", "process") == "inproc" else _mp


def start_server(capacity, batch_size, host=None, port=None, samplers=None, cache_policy=None, **kw):
    return _mod().start_server(capacity, batch_size, host, port, samplers, cache_policy, **kw)


def start_per(capacity, batch_size, alpha=0.6, beta=0.4, sample_start=1000, num_sampler_procs=1,
              host=None, port=None, cache_policy=None, **kw):
    return _mod().start_per(capacity, batch_size, alpha, beta, sample_start, num_sampler_procs, host, port,
                            cache_policy, **kw)
