#!/bin/bash
g++ -O3 -pthread -o /tmp/trim_probe tools/trim_probe.cpp || exit 1
for t in 1 8 12 16; do for m in 0 1; do /tmp/trim_probe $t $m | tail -2 | tr '\n' ' '; echo; done; done
