#!/bin/bash
for K in 64 128 32; do for s in 0 1 2 3 4 5 8 9 58; do for b in 0 1; do
PL_DBG_ASHIFT=$s PL_DBG_BASEOFF=$b timeout 60 python tools/dbg_rowshift.py $K 2>&1 | tail -1
done; done; done
