# two processes of the C++ facade test on two GPUs, bounded in time, outputs kept
set -u
D=/tmp/facade2; mkdir -p $D; rm -f $D/*
g++ -std=c++17 -O2 -ffp-contract=off -Iinclude tests/cpp/domain_facade_test.cpp -Lcornerstone_octree_b200 -lcstone_b200 -Wl,-rpath,$PWD/cornerstone_octree_b200 -o $D/exe || exit 1
for r in 0 1; do (timeout ${T:-90} $D/exe $r 2 $D/id 6000 > $D/out$r.log 2>&1; echo "rank $r exit $?" >> $D/out$r.log) & done
wait
for r in 0 1; do echo "== rank $r"; tail -${LINES_OUT:-14} $D/out$r.log; done
