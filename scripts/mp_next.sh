#!/bin/bash
RETH_APPEND_TILED=1 timeout 200 python scripts/append_ab.py 2>/dev/null | tail -1
RETH_APPEND_TILED=1 timeout 200 python scripts/append_ab.py 2>/dev/null | tail -1
