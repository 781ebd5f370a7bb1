import sys, types
sys.path.insert(0, "/root/repo")
import bench
for keys, bucket, n in (("u32", 16, 4000000), ("u64", 16, 4000000), ("u32", 64, 4000000), ("u32", 16, 16000000)):
    a = types.SimpleNamespace(n=n, dist="uniform", global_bucket=65536, bucket=bucket, ngmax=150, keys=keys, curve="hilbert")
    print(keys, bucket, n, bench.parity_in_run(a, n_check=n), flush=True)
