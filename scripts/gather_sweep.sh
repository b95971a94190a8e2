RETH_GATHER_MODE=1 python scripts/gather_sweep.py
RETH_GATHER_MODE=2 python scripts/gather_sweep.py
RETH_GATHER_MODE=2 RETH_BULK_EF_LOADS=0 python scripts/gather_sweep.py
RETH_GATHER_MODE=2 RETH_BULK_EF_STORES=1 python scripts/gather_sweep.py
RETH_GATHER_MODE=2 RETH_BULK_UNITS=882 python scripts/gather_sweep.py
RETH_GATHER_MODE=2 RETH_BULK_UNITS=882 RETH_BULK_CTAS=2 python scripts/gather_sweep.py
RETH_GATHER_MODE=2 RETH_BULK_CTAS=2 python scripts/gather_sweep.py
RETH_GATHER_MODE=2 RETH_BULK_UNITS=441 RETH_BULK_CTAS=4 python scripts/gather_sweep.py
