#!/bin/bash
for t in 0 1 0 1; do CAP=1048576 RETH_APPEND_TILED=$t timeout 200 python scripts/append_ab.py 2>/dev/null | tail -1 | sed "s/^/tiled=$t /"; done
