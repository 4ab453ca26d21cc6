#!/bin/bash
run() { python tools/long_probe.py 300 2>&1 | tail -1 | cut -c1-150; }
for i in 1 2 3; do run; done
