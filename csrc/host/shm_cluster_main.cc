// Stand-alone run of the shared-memory restatement of the device protocol (shm_cluster.h): every rank a thread,
// free-running rounds with random stalls, every output checked against its contributor counts.
//
//   aar_shm_cluster [world n maxChunkSize maxLag rounds thReduce thComplete thAllreduce stallProb stallUs timeoutMs
//                    skipNotices(1)]
//
// Exit code 1 when a sum does not match its count.  Built plain and under ThreadSanitizer by
// scripts/sanitize_host.sh: the data words are ordinary floats, so any read not ordered behind its writer by a
// flag shows up as a race report.
#include <cstdio>
#include <cstdlib>

#include "shm_cluster.h"

int main(int argc, char** argv) {
  auto arg = [&](int i, double dflt) { return argc > i ? std::atof(argv[i]) : dflt; };
  aar::ShmConfig cfg;
  cfg.world = static_cast<int>(arg(1, 4));
  cfg.n = static_cast<int64_t>(arg(2, 778));
  cfg.max_chunk = static_cast<int64_t>(arg(3, 3));
  cfg.max_lag = static_cast<int>(arg(4, 3));
  const int rounds = static_cast<int>(arg(5, 1000));
  cfg.th_reduce = static_cast<float>(arg(6, 1.0));
  cfg.th_complete = static_cast<float>(arg(7, 1.0));
  cfg.th_allreduce = static_cast<float>(arg(8, 1.0));
  const double stall_prob = arg(9, 0.0);
  const int stall_us = static_cast<int>(arg(10, 0));
  cfg.timeout_ms = arg(11, 2000.0);
  cfg.skip_notice = arg(12, 1.0) != 0.0;
  try {
    aar::ShmCluster cluster(cfg);
    const aar::ShmFreeRunStats s = cluster.run_free(rounds, stall_prob, stall_us, 1);
    const double elems = static_cast<double>(cfg.n) * cfg.world * rounds;
    std::printf(
        "finished: %d ranks, n %lld, maxChunkSize %lld, maxLag %d, th %.2f/%.2f/%.2f: %d rounds in %.3f s = %.1f rounds/s, "
        "%.4f MBytes/sec per rank (output check included); skip notices %s; partial %.2f%%, missing %.2f%%, %d timeouts, %d forced, %lld mismatches\n",
        cfg.world, static_cast<long long>(cfg.n), static_cast<long long>(cfg.max_chunk), cfg.max_lag, cfg.th_reduce,
        cfg.th_complete, cfg.th_allreduce, s.rounds, s.seconds, s.rounds / s.seconds,
        static_cast<double>(cfg.n) * 4 * s.rounds / 1e6 / s.seconds, cfg.skip_notice ? "on" : "off",
        100.0 * s.partial_elements / elems,
        100.0 * s.missing_elements / elems, s.timeouts, s.forced, static_cast<long long>(s.mismatches));
    return s.mismatches == 0 ? 0 : 1;
  } catch (const std::exception& e) {
    std::fprintf(stderr, "aar_shm_cluster: %s\n", e.what());
    return 2;
  }
}
