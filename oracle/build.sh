#!/bin/sh
# Builds the CPU oracle (test infrastructure) into oracle/libhs_oracle.so.
# The reference itself is Scala on Spark 2.4.2 and cannot be compiled in this image
# (no JVM, no sbt, no Spark jars), so there is no oracle/_ref.
set -e
cd "$(dirname "$0")"
gcc -O2 -fopenmp -fPIC -shared -std=c11 -Wall -o libhs_oracle.so hs_oracle.c -lm
