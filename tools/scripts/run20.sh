#!/bin/bash
cd /root/repo
CPN_MLP_TS=1 timeout 120 python tools/run_op.py mlp --M 46656 --C 384 --iters 1 2>&1 | tail -56
CPN_MLP_TS=1 CPN_MLP_CG=1 timeout 120 python tools/run_op.py mlp --M 50176 --C 192 --iters 1 2>&1 | tail -30
