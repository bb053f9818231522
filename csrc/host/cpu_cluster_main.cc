// aar_cpu_cluster — the host runtime without Python: one coordinator and N workers of the CPU backend in one
// process (threads, loopback TCP), with the native ramp source and throughput/assert sink.
//
//   aar_cpu_cluster [numOfWorkers=2] [dataSize=10] [maxChunkSize=2] [maxRound=1000] [maxLag=1]
//                   [thAllreduce=1] [thReduce=1] [thComplete=0.8]
//
// Same defaults as the reference's master launcher (AllreduceMaster.scala:98-107); exits 0 when every round has
// passed the master's gate and no sink assertion failed.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <vector>

#include "node.h"

using namespace aar;

int main(int argc, char** argv) {
  auto arg = [&](int i, double dflt) { return argc > i ? std::atof(argv[i]) : dflt; };
  MasterConfig cfg;
  cfg.total_workers = static_cast<int>(arg(1, 2));
  cfg.data_size = static_cast<int64_t>(arg(2, 10));
  cfg.max_chunk_size = static_cast<int64_t>(arg(3, 2));
  cfg.max_round = static_cast<int>(arg(4, 1000));
  cfg.max_lag = static_cast<int>(arg(5, 1));
  cfg.th_allreduce = static_cast<float>(arg(6, 1.0));
  cfg.th_reduce = static_cast<float>(arg(7, 1.0));
  cfg.th_complete = static_cast<float>(arg(8, 0.8));
  const bool exact = cfg.th_allreduce == 1.f && cfg.th_reduce == 1.f && cfg.th_complete == 1.f;

  MasterNode master(cfg);
  const int port = master.listen_on("127.0.0.1", 0);
  master.start();
  std::vector<std::unique_ptr<WorkerNode>> workers;
  std::vector<std::shared_ptr<ThroughputStats>> stats;
  const auto t0 = std::chrono::steady_clock::now();
  for (int i = 0; i < cfg.total_workers; ++i) {
    stats.push_back(std::make_shared<ThroughputStats>());
    workers.push_back(std::make_unique<WorkerNode>(
        PeerAddress{"127.0.0.1", port}, make_ramp_source(cfg.data_size),
        make_throughput_sink(cfg.data_size, cfg.max_round / 4 > 0 ? cfg.max_round / 4 : 1, exact ? cfg.total_workers : 0,
                             stats.back())));
    workers.back()->begin();
  }
  const bool ok = master.wait_finished(600.0);
  const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  for (auto& w : workers) w->wait_shutdown(5.0);
  int failures = 0;
  for (auto& s : stats) failures += s->failures.load();
  for (auto& w : workers) w->stop();
  master.stop();
  std::printf("%s: %d workers, dataSize %lld, maxChunkSize %lld: %d rounds in %.3f s = %.1f rounds/s, %.4f MBytes/sec per worker, "
              "%d sink assertion failures\n",
              ok ? "finished" : "TIMED OUT", cfg.total_workers, static_cast<long long>(cfg.data_size),
              static_cast<long long>(cfg.max_chunk_size), cfg.max_round + 1, secs, (cfg.max_round + 1) / secs,
              cfg.data_size * 4.0 * (cfg.max_round + 1) / 1e6 / secs, failures);
  return ok && failures == 0 ? 0 : 1;
}
