#!/bin/bash
cd /root/repo
CPN_MLP_TS=1 CPN_MLP_DBG=256 timeout 120 python tools/run_op.py mlp --M 46656 --C 384 --iters 1 2>&1 | tail -62
