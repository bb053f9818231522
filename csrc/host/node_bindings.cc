// pybind11 surface of the TCP nodes (coordinator + worker processes).
#include <pybind11/functional.h>
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include "node.h"

namespace py = pybind11;

namespace aar {

namespace {
using FloatArray = py::array_t<float, py::array::c_style | py::array::forcecast>;

DataSource py_source(py::object fn) {
  return [fn](int iteration) {
    py::gil_scoped_acquire gil;
    FloatArray a = fn(iteration).cast<FloatArray>();
    auto b = a.request();
    const float* p = static_cast<const float*>(b.ptr);
    return std::vector<float>(p, p + b.size);
  };
}
DataSink py_sink(py::object fn) {
  return [fn](const std::vector<float>& d, const std::vector<int>& c, int iteration) {
    py::gil_scoped_acquire gil;
    fn(py::array_t<float>(d.size(), d.data()), py::array_t<int>(c.size(), c.data()), iteration);
  };
}
}  // namespace

void bind_nodes(py::module_& m) {
  py::class_<MasterNodeOptions>(m, "MasterNodeOptions")
      .def(py::init<>())
      .def_readwrite("down_after_s", &MasterNodeOptions::down_after_s)
      .def_readwrite("shutdown_workers_on_finish", &MasterNodeOptions::shutdown_workers_on_finish)
      .def_readwrite("verbose", &MasterNodeOptions::verbose);

  py::class_<MasterNode>(m, "MasterNode")
      .def(py::init<MasterConfig, MasterNodeOptions>(), py::arg("config"),
           py::arg("options") = MasterNodeOptions{})
      .def("listen", &MasterNode::listen_on, py::arg("host") = "127.0.0.1", py::arg("port") = 0)
      .def("start", &MasterNode::start)
      .def("stop", &MasterNode::stop, py::call_guard<py::gil_scoped_release>())
      .def("wait_finished", &MasterNode::wait_finished, py::call_guard<py::gil_scoped_release>())
      .def_property_readonly("port", &MasterNode::port)
      .def_property_readonly("round", &MasterNode::round)
      .def_property_readonly("registered", &MasterNode::registered)
      .def_property_readonly("alive", &MasterNode::alive);

  py::class_<ThroughputStats, std::shared_ptr<ThroughputStats>>(m, "ThroughputStats")
      .def(py::init<>())
      .def_property_readonly("checkpoints", [](const ThroughputStats& s) { return s.checkpoints.load(); })
      .def_property_readonly("failures", [](const ThroughputStats& s) { return s.failures.load(); })
      .def_property_readonly("last_mbytes_per_s",
                             [](const ThroughputStats& s) { return s.last_mbytes_per_s.load(); });

  py::class_<WorkerNodeOptions>(m, "WorkerNodeOptions")
      .def(py::init<>())
      .def_readwrite("heartbeat_s", &WorkerNodeOptions::heartbeat_s)
      .def_readwrite("engine", &WorkerNodeOptions::engine);

  py::class_<WorkerNode>(m, "WorkerNode")
      .def(py::init([](const std::string& host, int port, py::object source, py::object sink,
                       WorkerNodeOptions opts) {
             DataSource src;
             DataSink snk;
             if (!source.is_none() && !sink.is_none()) {
               src = py_source(source);
               snk = py_sink(sink);
             }
             return std::make_unique<WorkerNode>(PeerAddress{host, port}, std::move(src), std::move(snk), opts);
           }),
           py::arg("master_host"), py::arg("master_port"), py::arg("source") = py::none(),
           py::arg("sink") = py::none(), py::arg("options") = WorkerNodeOptions{})
      .def_static(
          "builtin",
          [](const std::string& host, int port, int64_t data_size, int checkpoint, int assert_multiple,
             std::shared_ptr<ThroughputStats> stats, WorkerNodeOptions opts) {
            return std::make_unique<WorkerNode>(PeerAddress{host, port}, make_ramp_source(data_size),
                                                make_throughput_sink(data_size, checkpoint, assert_multiple, stats),
                                                opts);
          },
          py::arg("master_host"), py::arg("master_port"), py::arg("data_size"), py::arg("checkpoint") = 50,
          py::arg("assert_multiple") = 0, py::arg("stats") = nullptr, py::arg("options") = WorkerNodeOptions{},
          "Worker with the native ramp source and throughput/assert sink")
      .def("set_handler",
           [](WorkerNode& w, py::object fn) {
             w.set_handler([fn](const Message& msg) {
               py::gil_scoped_acquire gil;
               fn(msg);
             });
           })
      .def("listen", &WorkerNode::listen_on, py::arg("host") = "127.0.0.1", py::arg("port") = 0)
      .def("begin", &WorkerNode::begin, py::call_guard<py::gil_scoped_release>())
      .def("stop", &WorkerNode::stop, py::call_guard<py::gil_scoped_release>())
      .def("report_complete", &WorkerNode::report_complete)
      .def("wait_shutdown", &WorkerNode::wait_shutdown, py::call_guard<py::gil_scoped_release>())
      .def_property_readonly("port", &WorkerNode::port)
      .def_property_readonly("worker_id", &WorkerNode::worker_id)
      .def_property_readonly("rounds_completed", &WorkerNode::rounds_completed)
      .def_property_readonly("engine", [](WorkerNode& w) { return w.engine(); },
                             py::return_value_policy::reference_internal);
}

}  // namespace aar
