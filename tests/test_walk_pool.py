"""Stand-alone stress test of the host walk pool (beforeholiday_b200/csrc/walk_pool.h): compiled with g++ alone (no torch,
no CUDA), many back-to-back runs with varying part counts, checks that every part ran exactly once and nothing hangs."""
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent

SRC = r'''
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include "walk_pool.h"
int main() {
    for (unsigned workers : {0u, 1u, 3u, 7u}) {
        mt_host::WalkPool& pool = *new mt_host::WalkPool(workers);   // pools are leak-only (destructor deleted)
        std::vector<std::atomic<int>> hits(64);
        long long total = 0;
        for (int it = 0; it < 20000; it++) {
            const unsigned parts = 1 + (unsigned)(it * 2654435761u >> 7) % (workers + 4);   // also parts > workers + 1
            for (auto& h : hits) h.store(0);
            std::atomic<long long> sum{0};
            pool.run(parts, [&](unsigned p) { hits[p].fetch_add(1); sum.fetch_add(p + 1); });
            for (unsigned p = 0; p < 64; p++)
                if (hits[p].load() != (p < parts ? 1 : 0)) { std::printf("FAIL workers=%u it=%d part=%u hits=%d\n", workers, it, p, hits[p].load()); return 1; }
            if (sum.load() != (long long)parts * (parts + 1) / 2) { std::printf("FAIL sum\n"); return 1; }
            total += parts;
        }
        std::printf("workers=%u ok (%lld parts)\n", workers, total);
        std::fflush(stdout);
    }
    // the process-wide pool
    std::atomic<int> n{0};
    auto& g = mt_host::WalkPool::get();
    for (int it = 0; it < 1000; it++) g.run(g.workers() + 1, [&](unsigned) { n.fetch_add(1); });
    if (n.load() != 1000 * (int)(g.workers() + 1)) return 1;
    std::printf("global pool ok (%u workers)\n", g.workers());
    return 0;
}
'''


def test_walk_pool_stress(tmp_path):
    src = tmp_path / "wp.cpp"
    src.write_text(SRC)
    exe = tmp_path / "wp"
    subprocess.run(["g++", "-O2", "-std=c++17", "-pthread", f"-I{ROOT / 'beforeholiday_b200' / 'csrc'}", str(src), "-o", str(exe)],
                   check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True, timeout=120)
    sys.stdout.write(out.stdout)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "global pool ok" in out.stdout
