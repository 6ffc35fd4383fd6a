"""Host-to-device rate of cfd_upload (pinned host array -> staging area -> line layout) beside a plain 1-D copy of the same
bytes.  Usage: python tools/bench_upload.py [n]   (n x n control volumes, default 4096)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import cfd_b200


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    g = cfd_b200.CavityFVM(n, n, order="redblack")
    g.init_lid_driven(reynolds=1000.0)
    ni, nj = n + 2, n + 2
    host = torch.zeros((nj, ni), dtype=torch.float64, pin_memory=True)
    dev = torch.empty((nj, ni), dtype=torch.float64, device="cuda")
    nbytes = ni * nj * 8
    for name, fn in (("plain 1-D copy", lambda: dev.copy_(host, non_blocking=True)),
                     ("cfd_upload", lambda: g.upload_ptr("u", host.data_ptr(), ni)),
                     ("cfd_download", lambda: g.download_ptr("u", host.data_ptr(), ni))):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        K = 20
        for _ in range(K):
            fn()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / K
        print(f"{name:16s} {dt * 1e3:7.3f} ms per field of {nbytes / 1e6:.0f} MB = {nbytes / dt / 1e9:5.1f} GB/s", flush=True)
    g.close()


if __name__ == "__main__":
    main()
