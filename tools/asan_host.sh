#!/bin/sh
# Build the host library's sources with AddressSanitizer + UndefinedBehaviorSanitizer and run tools/asan_host_driver.cpp
# over two fleets (no GPU involved).  Takes about a minute; not part of the default test run.
#   sh tools/asan_host.sh          -> "asan driver done" and exit 0 when clean
set -e
root=$(cd "$(dirname "$0")/.." && pwd)
out="$root/build/asan"
mkdir -p "$out"
flags="-O1 -g -std=c++17 -pthread -fsanitize=address,undefined -fno-omit-frame-pointer"
for f in "$root"/odh-kubeflow_b200/csrc/host/*.cpp "$root"/tools/asan_host_driver.cpp; do
  g++ $flags -c "$f" -o "$out/$(basename "$f").o" &
done
wait
g++ $flags "$out"/*.o -o "$out/driver"
ASAN_OPTIONS=detect_leaks=1 "$out/driver"
