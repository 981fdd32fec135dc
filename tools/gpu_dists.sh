#!/bin/bash
D=${1:-gpu-radix-sort_b200/libsort/devbench}
$D check
for dist in uniform sorted consthigh zipf allsame; do $D time 67108864 5 $dist | grep -E "input|full32"; done
$D time 268435456 10 | grep full32
