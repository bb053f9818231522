// Python bindings of the shared-memory restatement of the device protocol (shm_cluster.h).
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include "shm_cluster.h"

namespace py = pybind11;

namespace aar {

using FloatArray = py::array_t<float, py::array::c_style | py::array::forcecast>;

static py::dict info_dict(const ShmRoundInfo& i) {
  py::dict d;
  d["reduced_full"] = i.reduced_full;
  d["reduced_partial"] = i.reduced_partial;
  d["chunks_present"] = i.chunks_present;
  d["timed_out"] = i.timed_out;
  d["dest_mask"] = i.dest_mask;
  return d;
}

void bind_shm(py::module_& m) {
  py::class_<ShmConfig>(m, "ShmConfig")
      .def(py::init<>())
      .def_readwrite("world", &ShmConfig::world)
      .def_readwrite("n", &ShmConfig::n)
      .def_readwrite("max_chunk", &ShmConfig::max_chunk)
      .def_readwrite("max_lag", &ShmConfig::max_lag)
      .def_readwrite("th_reduce", &ShmConfig::th_reduce)
      .def_readwrite("th_complete", &ShmConfig::th_complete)
      .def_readwrite("th_allreduce", &ShmConfig::th_allreduce)
      .def_readwrite("timeout_ms", &ShmConfig::timeout_ms)
      .def_readwrite("scale", &ShmConfig::scale)
      .def_readwrite("normalize", &ShmConfig::normalize)
      .def_readwrite("skip_notice", &ShmConfig::skip_notice);

  py::class_<ShmCluster>(m, "ShmCluster")
      .def(py::init<const ShmConfig&>())
      .def_static(
          "attach",
          [](const ShmConfig& cfg, const std::string& name, bool create, double timeout_s, bool accept_orphan) {
            py::gil_scoped_release nogil;
            return std::make_unique<ShmCluster>(cfg, name, create, timeout_s, accept_orphan);
          },
          py::arg("config"), py::arg("name"), py::arg("create"), py::arg("timeout_s") = 30.0,
          py::arg("accept_orphan") = false,
          "Cluster whose regions live in the named POSIX shared-memory segment (one process per rank)")
      .def(
          "rank_round",
          [](ShmCluster& c, int rank, int round, FloatArray src) {
            const int64_t n = c.layout().n;
            if (src.size() != n) throw std::invalid_argument("Input data size is different from initialization time!");
            py::array_t<float> out(n);
            py::array_t<int32_t> counts(n);
            ShmRoundInfo info;
            {
              const float* in = src.data();
              float* o = out.mutable_data();
              int32_t* k = counts.mutable_data();
              py::gil_scoped_release nogil;
              info = c.rank_round(rank, round, in, o, k, nullptr);
            }
            return py::make_tuple(out, counts, info_dict(info));
          },
          py::arg("rank"), py::arg("round"), py::arg("src"),
          "One rank's round (blocks until that rank has completed it): returns (out, counts, info)")
      .def_property_readonly("req_reduce", &ShmCluster::req_reduce)
      .def_property_readonly("req_complete", &ShmCluster::req_complete)
      .def_property_readonly("gate_count", &ShmCluster::gate_count)
      .def_property_readonly("total_chunks", &ShmCluster::total_chunks)
      .def_property_readonly("region_bytes", [](const ShmCluster& c) { return c.layout().total_bytes; })
      .def_property_readonly("next_round", &ShmCluster::next_round)
      .def("issued", &ShmCluster::issued)
      .def("force_complete", &ShmCluster::force_complete)
      .def("set_dead_peers", &ShmCluster::set_dead_peers, py::arg("rank"), py::arg("mask"))
      .def(
          "run_round",
          [](ShmCluster& c, int round, std::vector<FloatArray> srcs, uint32_t mute_mask, std::vector<uint32_t> skip_dest,
             std::vector<double> delay_us) {
            const size_t P = srcs.size();
            const int64_t n = c.layout().n;
            std::vector<const float*> in(P);
            std::vector<py::array_t<float>> outs;
            std::vector<py::array_t<int32_t>> counts;
            std::vector<float*> op(P);
            std::vector<int32_t*> cp(P);
            for (size_t r = 0; r < P; ++r) {
              if (srcs[r].size() != n) throw std::invalid_argument("Input data size is different from initialization time!");
              in[r] = srcs[r].data();
              outs.emplace_back(n);
              counts.emplace_back(n);
              op[r] = outs[r].mutable_data();
              cp[r] = counts[r].mutable_data();
            }
            ShmInjection inj;
            inj.mute_mask = mute_mask;
            for (size_t r = 0; r < skip_dest.size() && r < inj.skip_dest.size(); ++r) inj.skip_dest[r] = skip_dest[r];
            for (size_t r = 0; r < delay_us.size() && r < inj.delay_ns.size(); ++r)
              inj.delay_ns[r] = static_cast<uint64_t>(delay_us[r] * 1e3);
            std::vector<ShmRoundInfo> infos;
            {
              py::gil_scoped_release nogil;
              infos = c.run_round(round, in, op, cp, &inj);
            }
            py::list il;
            for (const auto& i : infos) il.append(info_dict(i));
            return py::make_tuple(outs, counts, il);
          },
          py::arg("round"), py::arg("srcs"), py::arg("mute_mask") = 0u, py::arg("skip_dest") = std::vector<uint32_t>{},
          py::arg("delay_us") = std::vector<double>{},
          "Round `round` on every rank (one thread each): returns (outs, counts, per-rank info)")
      .def(
          "run_free",
          [](ShmCluster& c, int rounds, double stall_prob, int stall_us, uint64_t seed) {
            ShmFreeRunStats s;
            {
              py::gil_scoped_release nogil;
              s = c.run_free(rounds, stall_prob, stall_us, seed);
            }
            py::dict d;
            d["rounds"] = s.rounds;
            d["seconds"] = s.seconds;
            d["rounds_per_s"] = s.seconds > 0 ? s.rounds / s.seconds : 0.0;
            d["mismatches"] = s.mismatches;
            d["partial_elements"] = s.partial_elements;
            d["missing_elements"] = s.missing_elements;
            d["timeouts"] = s.timeouts;
            d["forced"] = s.forced;
            return d;
          },
          py::arg("rounds"), py::arg("stall_prob") = 0.0, py::arg("stall_us") = 0, py::arg("seed") = 1,
          "Every rank free-runs `rounds` rounds with random stalls; sums are checked against counts");
}

}  // namespace aar
