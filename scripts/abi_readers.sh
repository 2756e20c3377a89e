#!/bin/bash
# Build and run tests/abi_readers.c (reader threads through the C ABI); extra environment is passed through.
set -e
cd "$(dirname "$0")/.."
L=modular-rag-system_b200/csrc
gcc -O2 -std=c11 -D_POSIX_C_SOURCE=200809L -I include tests/abi_readers.c -o /tmp/abi_readers $L/libragb200.so -lpthread -lm -Wl,-rpath,$PWD/$L
/tmp/abi_readers "${1:-1.0}"
