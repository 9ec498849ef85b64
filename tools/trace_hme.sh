#!/bin/bash
# Builds the tracing variant of the library on the GPU box (ME_HME_TRACE: per-CTA globaltimer stamps of every stage),
# runs a few pictures of the bench workload and prints the trace lines. The box copy is scratch, nothing to restore.
set -e
( cd svt-av1-psyex_b200 && make -B EXTRA=-DME_HME_TRACE >/dev/null 2>&1 )
python tools/profile_step.py ${1:-2} 2>&1 | grep -E "^b |^poc|^CTA"
