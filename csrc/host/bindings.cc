// pybind11 surface of the host library (protocol math, buffers, engines, bus, TCP nodes).
#include <pybind11/functional.h>
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <memory>

#include "aar/protocol_math.h"
#include "buffers.h"
#include "local_bus.h"
#include "master_engine.h"
#include "messages.h"
#include "node.h"
#include "worker_engine.h"

namespace aar {
void bind_shm(pybind11::module_& m);
}

namespace py = pybind11;
using namespace aar;

namespace {

using FloatArray = py::array_t<float, py::array::c_style | py::array::forcecast>;

std::vector<float> to_vec(const FloatArray& a) {
  auto b = a.request();
  const float* p = static_cast<const float*>(b.ptr);
  return std::vector<float>(p, p + b.size);
}
py::array_t<float> to_np(const std::vector<float>& v) { return py::array_t<float>(v.size(), v.data()); }
py::array_t<int> to_np(const std::vector<int>& v) { return py::array_t<int>(v.size(), v.data()); }

DataSource wrap_source(py::object fn) {
  return [fn](int iteration) {
    py::gil_scoped_acquire gil;
    py::object out = fn(iteration);
    return to_vec(out.cast<FloatArray>());
  };
}
DataSink wrap_sink(py::object fn) {
  return [fn](const std::vector<float>& d, const std::vector<int>& c, int iteration) {
    py::gil_scoped_acquire gil;
    fn(to_np(d), to_np(c), iteration);
  };
}

// A LocalBus that also owns the engines plugged into it.
struct PyBus {
  LocalBus bus;
  std::vector<std::unique_ptr<WorkerEngine>> workers;
  std::vector<std::unique_ptr<MasterEngine>> masters;
};

}  // namespace

PYBIND11_MODULE(_aar_host, m) {
  m.doc() = "Host library of akka_allreduce_b200: protocol engines, buffers, bus, TCP nodes";

  m.def("threshold_count", &threshold_count, "int(float32(th)*float32(n))");
  m.def("master_gate", &master_gate);
  m.def("num_chunks", &num_chunks);
  m.def("total_reduced_chunks_ref", &total_reduced_chunks_ref);

  py::class_<Partition>(m, "Partition")
      .def(py::init<int64_t, int>())
      .def_readonly("n", &Partition::n)
      .def_readonly("P", &Partition::P)
      .def_readonly("step", &Partition::step)
      .def("start", &Partition::start)
      .def("end", &Partition::end)
      .def("size", &Partition::size)
      .def("max_block", &Partition::max_block)
      .def("min_block", &Partition::min_block)
      .def("reference_defined", &Partition::reference_defined)
      .def("total_reduced_chunks", [](const Partition& p, int64_t mc) { return total_reduced_chunks(p, mc); });

  // ---- messages -------------------------------------------------------------------
  py::class_<InitWorkers>(m, "InitWorkers")
      .def(py::init([](std::map<int, Endpoint> workers, int worker_num, Endpoint master, int dest_id,
                       float th_reduce, float th_complete, int max_lag, int64_t data_size,
                       int64_t max_chunk_size, int start_round) {
             InitWorkers x;
             x.workers = std::move(workers);
             x.worker_num = worker_num;
             x.master = master;
             x.dest_id = dest_id;
             x.th_reduce = th_reduce;
             x.th_complete = th_complete;
             x.max_lag = max_lag;
             x.data_size = data_size;
             x.max_chunk_size = max_chunk_size;
             x.start_round = start_round;
             return x;
           }),
           py::arg("workers"), py::arg("worker_num"), py::arg("master"), py::arg("dest_id"),
           py::arg("th_reduce"), py::arg("th_complete"), py::arg("max_lag"), py::arg("data_size"),
           py::arg("max_chunk_size"), py::arg("start_round") = 0)
      .def_readwrite("workers", &InitWorkers::workers)
      .def_readwrite("worker_num", &InitWorkers::worker_num)
      .def_readwrite("master", &InitWorkers::master)
      .def_readwrite("dest_id", &InitWorkers::dest_id)
      .def_readwrite("th_reduce", &InitWorkers::th_reduce)
      .def_readwrite("th_complete", &InitWorkers::th_complete)
      .def_readwrite("max_lag", &InitWorkers::max_lag)
      .def_readwrite("data_size", &InitWorkers::data_size)
      .def_readwrite("max_chunk_size", &InitWorkers::max_chunk_size)
      .def_readwrite("start_round", &InitWorkers::start_round)
      .def_readwrite("max_round", &InitWorkers::max_round)
      .def("__repr__", [](const InitWorkers& x) { return describe(Message(x)); });

  py::class_<StartAllreduce>(m, "StartAllreduce")
      .def(py::init([](int round) { return StartAllreduce{round}; }), py::arg("round"))
      .def_readwrite("round", &StartAllreduce::round)
      .def("__eq__", [](const StartAllreduce& a, const StartAllreduce& b) { return a.round == b.round; })
      .def("__repr__", [](const StartAllreduce& x) { return describe(Message(x)); });

  py::class_<ScatterBlock>(m, "ScatterBlock")
      .def(py::init([](FloatArray value, int src_id, int dest_id, int chunk_id, int round) {
             ScatterBlock x;
             x.value = to_vec(value);
             x.src_id = src_id;
             x.dest_id = dest_id;
             x.chunk_id = chunk_id;
             x.round = round;
             return x;
           }),
           py::arg("value"), py::arg("src_id"), py::arg("dest_id"), py::arg("chunk_id"), py::arg("round"))
      .def_property_readonly("value", [](const ScatterBlock& x) { return to_np(x.value); })
      .def_readwrite("src_id", &ScatterBlock::src_id)
      .def_readwrite("dest_id", &ScatterBlock::dest_id)
      .def_readwrite("chunk_id", &ScatterBlock::chunk_id)
      .def_readwrite("round", &ScatterBlock::round)
      .def("__eq__",
           [](const ScatterBlock& a, const ScatterBlock& b) {
             return a.value == b.value && a.src_id == b.src_id && a.dest_id == b.dest_id &&
                    a.chunk_id == b.chunk_id && a.round == b.round;
           })
      .def("__repr__", [](const ScatterBlock& x) { return describe(Message(x)); });

  py::class_<ReduceBlock>(m, "ReduceBlock")
      .def(py::init([](FloatArray value, int src_id, int dest_id, int chunk_id, int round, int count) {
             ReduceBlock x;
             x.value = to_vec(value);
             x.src_id = src_id;
             x.dest_id = dest_id;
             x.chunk_id = chunk_id;
             x.round = round;
             x.count = count;
             return x;
           }),
           py::arg("value"), py::arg("src_id"), py::arg("dest_id"), py::arg("chunk_id"), py::arg("round"),
           py::arg("count"))
      .def_property_readonly("value", [](const ReduceBlock& x) { return to_np(x.value); })
      .def_readwrite("src_id", &ReduceBlock::src_id)
      .def_readwrite("dest_id", &ReduceBlock::dest_id)
      .def_readwrite("chunk_id", &ReduceBlock::chunk_id)
      .def_readwrite("round", &ReduceBlock::round)
      .def_readwrite("count", &ReduceBlock::count)
      .def("__eq__",
           [](const ReduceBlock& a, const ReduceBlock& b) {
             return a.value == b.value && a.src_id == b.src_id && a.dest_id == b.dest_id &&
                    a.chunk_id == b.chunk_id && a.round == b.round && a.count == b.count;
           })
      .def("__repr__", [](const ReduceBlock& x) { return describe(Message(x)); });

  py::class_<CompleteAllreduce>(m, "CompleteAllreduce")
      .def(py::init([](int src_id, int round) { return CompleteAllreduce{src_id, round}; }),
           py::arg("src_id"), py::arg("round"))
      .def_readwrite("src_id", &CompleteAllreduce::src_id)
      .def_readwrite("round", &CompleteAllreduce::round)
      .def("__eq__", [](const CompleteAllreduce& a,
                        const CompleteAllreduce& b) { return a.src_id == b.src_id && a.round == b.round; })
      .def("__repr__", [](const CompleteAllreduce& x) { return describe(Message(x)); });

  // ---- buffers --------------------------------------------------------------------
  py::class_<ScatterBuffer>(m, "ScatterBuffer")
      .def(py::init<int64_t, int, int, float, int64_t>(), py::arg("block_size"), py::arg("peers"),
           py::arg("slots"), py::arg("th_reduce"), py::arg("max_chunk"))
      .def("store", [](ScatterBuffer& b, FloatArray d, int row, int src,
                       int chunk) { b.store(d.data(), static_cast<size_t>(d.size()), row, src, chunk); })
      .def("advance", &ScatterBuffer::advance)
      .def("at_threshold", &ScatterBuffer::at_threshold)
      .def("arrivals", &ScatterBuffer::arrivals)
      .def("reduce",
           [](const ScatterBuffer& b, int row, int chunk) {
             auto [s, c] = b.reduce(row, chunk);
             return py::make_tuple(to_np(s), c);
           })
      .def_property_readonly("required", &ScatterBuffer::required)
      .def_property_readonly("chunks", &ScatterBuffer::chunks)
      .def_property_readonly("slots", &ScatterBuffer::slots)
      .def_property_readonly("peers", &ScatterBuffer::peers)
      .def_property_readonly("width", &ScatterBuffer::width);

  py::class_<ReduceBuffer>(m, "ReduceBuffer")
      .def(py::init<int64_t, int64_t, int64_t, int, int, float, int64_t>(), py::arg("max_block"),
           py::arg("min_block"), py::arg("total_size"), py::arg("peers"), py::arg("slots"),
           py::arg("th_complete"), py::arg("max_chunk"))
      .def("store", [](ReduceBuffer& b, FloatArray d, int row, int src, int chunk,
                       int count) { b.store(d.data(), static_cast<size_t>(d.size()), row, src, chunk, count); })
      .def("advance", &ReduceBuffer::advance)
      .def("at_threshold", &ReduceBuffer::at_threshold)
      .def("arrivals", &ReduceBuffer::arrivals)
      .def("assemble",
           [](const ReduceBuffer& b, int row) {
             auto [d, c] = b.assemble(row);
             return py::make_tuple(to_np(d), to_np(c));
           })
      .def_property_readonly("required", &ReduceBuffer::required)
      .def_property_readonly("chunks", &ReduceBuffer::chunks)
      .def_property_readonly("slots", &ReduceBuffer::slots)
      .def_property_readonly("peers", &ReduceBuffer::peers)
      .def_property_readonly("width", &ReduceBuffer::width);

  // ---- engines + bus --------------------------------------------------------------
  py::class_<WorkerOptions>(m, "WorkerOptions")
      .def(py::init<>())
      .def_readwrite("rethrow", &WorkerOptions::rethrow)
      .def_readwrite("strict_partition", &WorkerOptions::strict_partition)
      .def_readwrite("max_round_jump", &WorkerOptions::max_round_jump)
      .def_readwrite("verbose", &WorkerOptions::verbose);

  py::class_<WorkerEngine>(m, "WorkerEngine")
      .def_property_readonly("id", &WorkerEngine::id)
      .def_property_readonly("round", &WorkerEngine::round)
      .def_property_readonly("max_round", &WorkerEngine::max_round)
      .def_property_readonly("max_scattered", &WorkerEngine::max_scattered)
      .def_property_readonly("peer_count", &WorkerEngine::peer_count)
      .def_property_readonly("completed", &WorkerEngine::completed)
      .def_property_readonly("errors", &WorkerEngine::errors)
      .def_property_readonly("scatter_required", &WorkerEngine::scatter_required)
      .def_property_readonly("complete_required", &WorkerEngine::complete_required);

  py::class_<MasterConfig>(m, "MasterConfig")
      .def(py::init<>())
      .def_readwrite("total_workers", &MasterConfig::total_workers)
      .def_readwrite("th_allreduce", &MasterConfig::th_allreduce)
      .def_readwrite("th_reduce", &MasterConfig::th_reduce)
      .def_readwrite("th_complete", &MasterConfig::th_complete)
      .def_readwrite("max_lag", &MasterConfig::max_lag)
      .def_readwrite("data_size", &MasterConfig::data_size)
      .def_readwrite("max_round", &MasterConfig::max_round)
      .def_readwrite("max_chunk_size", &MasterConfig::max_chunk_size)
      .def_readwrite("announce_removals", &MasterConfig::announce_removals);

  py::class_<MasterEngine>(m, "MasterEngine")
      .def("member_up", &MasterEngine::member_up)
      .def("terminated", &MasterEngine::terminated)
      .def_property_readonly("round", &MasterEngine::round)
      .def_property_readonly("num_complete", &MasterEngine::num_complete)
      .def_property_readonly("workers", &MasterEngine::workers)
      .def_property_readonly("finished", &MasterEngine::finished);

  py::class_<PyBus>(m, "LocalBus")
      .def(py::init<>())
      .def("add_probe", [](PyBus& b) { return b.bus.add_probe(); })
      .def(
          "add_worker",
          [](py::object self, py::object source, py::object sink, WorkerOptions opts) {
            PyBus& b = self.cast<PyBus&>();
            const Endpoint ep = b.bus.add_endpoint(nullptr);
            b.workers.push_back(
                std::make_unique<WorkerEngine>(ep, wrap_source(source), wrap_sink(sink), &b.bus, opts));
            WorkerEngine* w = b.workers.back().get();
            b.bus.rebind(ep, [w](const Message& msg) { w->handle(msg); });
            // the bus owns the engine (and is its mailer): the handle keeps the bus alive, not the other way round
            return py::make_tuple(ep, py::cast(w, py::return_value_policy::reference_internal, self));
          },
          py::arg("source"), py::arg("sink"), py::arg("options") = WorkerOptions{})
      .def(
          "add_master",
          [](py::object self, MasterConfig cfg, bool verbose) {
            PyBus& b = self.cast<PyBus&>();
            const Endpoint ep = b.bus.add_endpoint(nullptr);
            b.masters.push_back(std::make_unique<MasterEngine>(ep, cfg, &b.bus, verbose));
            MasterEngine* me = b.masters.back().get();
            b.bus.rebind(ep, [me](const Message& msg) { me->handle(msg); });
            return py::make_tuple(ep, py::cast(me, py::return_value_policy::reference_internal, self));
          },
          py::arg("config"), py::arg("verbose") = false)
      .def("send", [](PyBus& b, Endpoint to, Message msg) { b.bus.send(to, std::move(msg)); })
      .def("run", [](PyBus& b, size_t max_steps) { return b.bus.run(max_steps); },
           py::arg("max_steps") = size_t(1) << 22)
      .def("set_chaos", [](PyBus& b, uint64_t seed, double drop, double delay) { b.bus.set_chaos(seed, drop, delay); },
           py::arg("seed"), py::arg("drop") = 0.0, py::arg("delay") = 0.0,
           "Drop / reorder data messages in transit (deterministic fault injection for tests)")
      .def("chaos_dropped", [](PyBus& b) { return b.bus.chaos_dropped(); })
      .def("pending", [](PyBus& b) { return b.bus.pending(); })
      .def("dropped", [](PyBus& b) { return b.bus.dropped(); })
      .def("take",
           [](PyBus& b, Endpoint probe) {
             std::vector<Message> out;
             out.swap(b.bus.inbox(probe));
             return out;
           },
           "Remove and return everything the probe has received so far");

  bind_nodes(m);
  bind_shm(m);
}
