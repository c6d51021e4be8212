#!/bin/bash
for k in 1 2 4 6 8; do SSPP_HOST_CHUNKS=$k python tools/e2e_probe.py 2>&1 | tail -1; done
for k in 4 6 8; do SSPP_HOST_CHUNKS=$k SSPP_HOST_CHUNK_SHAPE=equal python tools/e2e_probe.py 2>&1 | tail -1; done
