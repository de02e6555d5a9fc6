#!/bin/bash
# usage: tests/run_ranks.sh SECONDS NPROC PORT LOGFILE script [args...]
# Launches one process per GPU (torchrun, rendezvous on 127.0.0.1) in its own process group and kills the WHOLE group
# after SECONDS -- `timeout torchrun ...` would leave hung workers behind, holding the caller's pipes open.
secs=$1; n=$2; port=$3; log=$4; shift 4
setsid python -m torch.distributed.run --nproc-per-node "$n" --master-addr 127.0.0.1 --master-port "$port" "$@" > "$log" 2>&1 < /dev/null &
pid=$!
# the watcher must not inherit the caller's stdout / stderr: its `sleep` outlives this script and would keep a pipe open
( sleep "$secs"; echo "run_ranks: limit of ${secs}s hit, killing group $pid" >> "$log"; kill -KILL -- "-$pid" 2>/dev/null ) > /dev/null 2>&1 < /dev/null &
watcher=$!
wait "$pid"; rc=$?
pkill -P "$watcher" 2>/dev/null; kill "$watcher" 2>/dev/null
echo "run_ranks: rc=$rc ($*)"
tail -n 4 "$log" | cut -c1-400
exit $rc
