#!/bin/bash
# strict-check.sh -- the two acceptance checks of the reference (bin/nbody-sanity-check and
# bin/tests.sh), rewritten with a verdict that can actually fail: the reference's tests.sh prints
# the diff but never writes it to $DIFF_FILE, so it stamps every run PASSED, and its
# failed-tests.sh greps a spelling tests.sh never writes.  Run from a
# directory that holds nbody/nbody-par, nbody.ppm and the two golden files.
#   usage: bin/strict-check.sh <golden-32x100000> <golden-128x1000> [extra nbody-par args...]
set -u
G32="$1"; G128="$2"; shift 2
here="$(cd "$(dirname "$0")" && pwd)"
PATH="$here:$PATH"
fail=0
tmp="$(mktemp -d)"
run() {  # np cpus n steps golden
    local out="$tmp/$1_$2_$3.out" err="$tmp/$1_$2_$3.err"
    prun -v -"$2" -np "$1" -sge-script none nbody/nbody-par "$3" 0 nbody.ppm "$4" "${@:6}" > "$out" 2> "$err"
    if ! grep -q "took" "$err"; then echo "STATUS=FAILED np=$1 cpus=$2 n=$3: no run time on stderr"; fail=1; return; fi
    if diff -q "$5" "$out" > /dev/null; then echo "STATUS=PASSED np=$1 cpus=$2 n=$3 steps=$4"
    else echo "STATUS=FAILED np=$1 cpus=$2 n=$3 steps=$4"; diff "$5" "$out" | head -5; fail=1; fi
}
run 2 1 32 100000 "$G32" "$@"
for np in 1 2 3 4 5 6 7 8; do
    for cpus in 1 4 16; do
        run "$np" "$cpus" 128 1000 "$G128" "$@"
    done
done
rm -rf "$tmp"
exit $fail
