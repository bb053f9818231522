"""Config file, CLI argument handling, and the TCP coordinator + CPU workers (multi-process plumbing)."""
import os
import subprocess
import sys
import time

import numpy as np
import pytest

from akka_allreduce_b200 import AllReduceInput, AllreduceConfig, load_config
from akka_allreduce_b200.config import parse_hocon, DEFAULT_CONFIG_PATH
from akka_allreduce_b200.master import AllreduceMaster, build_config
from akka_allreduce_b200.worker import AllreduceWorker

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_hocon_subset():
    t = parse_hocon('a { b = 1, c: "x y" }\n a.d = 2.5 # c\n e = [1, 2]\n f { g { h = 10s } }\n// c2\n i = off')
    assert t == {"a": {"b": 1, "c": "x y", "d": 2.5}, "e": [1, 2], "f": {"g": {"h": 10.0}}, "i": False}


def test_hocon_parser_rejects_garbage_with_valueerror_only():
    """Any text either parses or raises ValueError (a config typo must not surface as an IndexError)."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=400, deadline=None)
    @given(st.text(alphabet='ab.= :{}[],"#/\n 19smh-$', max_size=24))
    def run(text):
        try:
            assert isinstance(parse_hocon(text), dict)
        except ValueError:
            pass

    run()
    for broken in ("a =", "a = [1, 2", "a { b = 1", "}", "a = ]"):
        with pytest.raises(ValueError):
            parse_hocon(broken)


def test_default_config_file_matches_reference_hardcoded_values():
    """AllreduceMaster.scala:98-107: maxLag 1, maxRound 100, th 1/1/0.8, 2 workers, dataSize 10, chunk 2."""
    c = load_config(DEFAULT_CONFIG_PATH)
    assert (c.thresholds.thAllreduce, c.thresholds.thReduce, c.thresholds.thComplete) == (1.0, 1.0, 0.8)
    assert (c.worker.totalSize, c.worker.maxLag, c.data.dataSize, c.data.maxChunkSize, c.data.maxRound) == (2, 1, 10, 2, 100)
    assert c.cluster.port == 2551 and c.cluster.downAfter == 10.0


def test_packaged_config_copy_is_identical():
    """The reference ships two byte-identical copies of its config (src/main/resources and conf/); so do we."""
    from akka_allreduce_b200.config import PACKAGED_CONFIG_PATH

    assert PACKAGED_CONFIG_PATH.read_bytes() == DEFAULT_CONFIG_PATH.read_bytes()


def test_reference_akka_conf_is_accepted():
    """The reference's own application.conf files load UNMODIFIED (nested blocks, dotted keys after the block,
    `${user.dir}` substitution, durations): master address from the first seed node, loglevel, auto-down timeout."""
    for ref in ("/root/reference/src/main/resources/application.conf", "/root/reference/conf/application.conf"):
        if not os.path.exists(ref):
            pytest.skip("reference tree not mounted")
        c = load_config(ref)
        assert (c.cluster.host, c.cluster.port) == ("127.0.0.1", 2551)
        assert c.cluster.downAfter == 10.0 and c.cluster.loglevel == "INFO"
        tree = parse_hocon(open(ref).read())
        assert tree["akka"]["cluster"]["metrics"]["enabled"] is False
        assert tree["akka"]["cluster"]["metrics"]["native-library-extract-folder"] == os.getcwd() + "/target/native"
        assert tree["akka"]["extensions"] == ["akka.cluster.metrics.ClusterMetricsExtension"]


def test_hocon_substitutions(monkeypatch):
    monkeypatch.setenv("AAR_TEST_VALUE", "42")
    t = parse_hocon("a = ${AAR_TEST_VALUE}\nb = ${AAR_TEST_VALUE}/x\nc = ${?AAR_NOT_SET_ANYWHERE}\nd { e = ${user.home} }")
    assert t["a"] == 42 and t["b"] == "42/x" and t["c"] == "${?AAR_NOT_SET_ANYWHERE}" and t["d"]["e"] == os.path.expanduser("~")


def test_master_cli_positional_order():
    """README.md:3 — PORT numOfWorkers dataSize maxChunkSize; dataSize defaults to 5*numOfWorkers."""
    c = build_config(["2600", "4", "778", "3", "--maxLag", "3", "--thComplete", "1.0"])
    assert (c.cluster.port, c.worker.totalSize, c.data.dataSize, c.data.maxChunkSize) == (2600, 4, 778, 3)
    assert c.worker.maxLag == 3 and c.thresholds.thComplete == 1.0
    assert build_config(["2600", "6"]).data.dataSize == 30
    assert build_config([]).cluster.port == 2551


def _cluster(P, n, mc, rounds, th=(1.0, 1.0, 1.0), max_lag=1):
    cfg = AllreduceConfig()
    cfg.cluster.port = 0
    cfg.worker.totalSize, cfg.worker.maxLag = P, max_lag
    cfg.data.dataSize, cfg.data.maxChunkSize, cfg.data.maxRound = n, mc, rounds
    cfg.thresholds.thAllreduce, cfg.thresholds.thReduce, cfg.thresholds.thComplete = th
    return cfg


def test_tcp_cluster_python_callbacks():
    """3 workers with Python source/sink over real sockets; thresholds 1 -> exact sums."""
    P, n = 3, 20
    cfg = _cluster(P, n, 3, 12, max_lag=2)
    master = AllreduceMaster(cfg)
    port = master.start()
    outs = [dict() for _ in range(P)]
    workers = []
    for i in range(P):
        def source(req):
            return AllReduceInput(np.arange(n, dtype=np.float32) + req.iteration)

        def sink(out, i=i):
            outs[i][out.iteration] = (np.array(out.data), np.array(out.count))

        w = AllreduceWorker(source, sink, master_port=port)
        w.start()
        workers.append(w)
    try:
        assert master.wait(30.0), f"master stuck at round {master.round}"
        assert all(w.wait(10.0) for w in workers)
    finally:
        for w in workers:
            w.stop()
        master.stop()
    assert sorted(w.id for w in workers) == [0, 1, 2]
    for o in outs:
        assert sorted(o) == list(range(13))
        for it, (d, c) in o.items():
            assert (d == P * (np.arange(n, dtype=np.float32) + it)).all() and (c == P).all()


def test_coordinator_survives_garbage_on_its_port():
    """Something that is not a worker connects to the coordinator port: an absurd frame length gets the connection
    closed, undecodable frames are dropped, and a normal job on the same coordinator still runs to the end."""
    import socket
    import struct

    P, n = 2, 12
    cfg = _cluster(P, n, 4, 10)
    master = AllreduceMaster(cfg)
    port = master.start()
    junk = []
    try:
        s1 = socket.create_connection(("127.0.0.1", port), timeout=5)
        s1.sendall(struct.pack("<IB", 0xFFFFFFF0, 2) + b"x" * 64)          # 4 GB frame announced
        s2 = socket.create_connection(("127.0.0.1", port), timeout=5)
        for kind in range(0, 12):                                             # every frame kind, truncated payloads
            s2.sendall(struct.pack("<IB", 3, kind) + b"\x01\x02\x03")
        s2.sendall(os.urandom(4096))
        junk = [s1, s2]
        s1.settimeout(5)
        assert s1.recv(1) == b""                                              # the coordinator hung up on s1
        outs = [dict() for _ in range(P)]
        workers = []
        for i in range(P):
            def source(req):
                return AllReduceInput(np.ones(n, dtype=np.float32))

            def sink(out, i=i):
                outs[i][out.iteration] = np.array(out.data)

            w = AllreduceWorker(source, sink, master_port=port)
            w.start()
            workers.append(w)
        assert master.wait(30.0), f"master stuck at round {master.round}"
        for w in workers:
            w.stop()
    finally:
        for s in junk:
            s.close()
        master.stop()
    assert all(sorted(o) == list(range(11)) and all((d == P).all() for d in o.values()) for o in outs)


def test_tcp_cluster_survives_dead_worker_with_thresholds():
    """Failure path (AllreduceMaster.scala:46-52): one of 4 workers dies; thresholds 0.75 keep rounds going
    and the master drops it from the set."""
    P, n = 4, 16
    cfg = _cluster(P, n, 2, 60, th=(0.75, 0.75, 0.75), max_lag=3)
    master = AllreduceMaster(cfg)
    port = master.start()
    done = [0] * P
    workers = []
    for i in range(P):
        def source(req):
            return AllReduceInput(np.ones(n, dtype=np.float32))

        def sink(out, i=i):
            done[i] = max(done[i], out.iteration)
            got = np.array(out.count) > 0
            assert (np.array(out.data)[got] == np.array(out.count)[got]).all()  # sum of ones == count

        w = AllreduceWorker(source, sink, master_port=port)
        w.start()
        workers.append(w)
    try:
        t0 = time.time()
        while master.round < 5 and time.time() - t0 < 20:
            time.sleep(0.01)
        workers[3].stop()  # kill one
        assert master.wait(30.0), f"master stuck at round {master.round}"
        assert master.alive == 3
    finally:
        for w in workers[:3]:
            w.stop()
        master.stop()
    assert all(d >= 55 for d in done[:3])


def test_cli_processes_default_plumbing_config():
    """BASELINE config 1: master + 2 worker processes via the CLIs, dataSize 10, chunk 2 (CPU only)."""
    env = dict(os.environ, PYTHONPATH=ROOT)
    port = 27000 + os.getpid() % 2000
    m = subprocess.Popen([sys.executable, "-m", "akka_allreduce_b200.master", str(port), "2", "10", "2",
                          "--maxRound", "200", "--thComplete", "1.0"], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    ws = [subprocess.Popen([sys.executable, "-m", "akka_allreduce_b200.worker", "0", "10", "--master-port", str(port),
                            "--checkpoint", "50", "--assert-multiple", "2"], env=env, stdout=subprocess.PIPE,
                           stderr=subprocess.STDOUT, text=True) for _ in range(2)]
    try:
        outs = [w.communicate(timeout=60)[0] for w in ws]
        mout = m.communicate(timeout=30)[0]
    finally:
        for p in ws + [m]:
            if p.poll() is None:
                p.kill()
    assert m.returncode == 0, mout
    for w, o in zip(ws, outs):
        assert w.returncode == 0, o
        assert "MBytes/sec" in o and "----Data output at #50" in o


def test_tcp_replacement_worker_rejoins():
    """Same elastic path over real sockets: kill one of 4 workers, start a replacement process-side worker,
    the master hands it the vacated id and every peer re-routes to its new address."""
    P, n = 4, 16
    cfg = _cluster(P, n, 2, 400, th=(0.75, 0.75, 0.75), max_lag=3)
    master = AllreduceMaster(cfg)
    port = master.start()
    seen = {}

    def make(i):
        def source(req):
            return AllReduceInput(np.ones(n, dtype=np.float32))

        def sink(out, i=i):
            seen.setdefault(i, []).append((out.iteration, np.array(out.count)))

        w = AllreduceWorker(source, sink, master_port=port)
        w.start()
        return w

    workers = [make(i) for i in range(P)]
    try:
        t0 = time.time()
        while master.round < 20 and time.time() - t0 < 20:
            time.sleep(0.005)
        victim = workers[3].id     # ids follow the order the joins REACHED the master, not list order
        workers[3].stop()
        t0 = time.time()
        while master.alive != 3 and time.time() - t0 < 10:
            time.sleep(0.01)
        assert master.alive == 3
        repl = make("replacement")
        workers.append(repl)
        assert master.wait(60), f"master stuck at round {master.round}"
        assert master.alive == 4 and repl.id == victim
    finally:
        for w in workers[:3] + workers[4:]:
            w.stop()
        master.stop()
    rounds = [r for r, _ in seen["replacement"]]
    assert rounds and min(rounds) >= 20          # joined at the current round
    survivor = next(i for i in range(3) if seen.get(i))
    lo = victim * (n // P)                       # the vacated block is being reduced again
    assert any(int(c[lo:lo + n // P].min()) > 0 for r, c in seen[survivor] if r > min(rounds) + 5)


def test_quickstart_example_runs():
    p = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "quickstart_cpu.py")], capture_output=True,
                       text=True, timeout=120, env=dict(os.environ, PYTHONPATH=ROOT))
    assert p.returncode == 0 and p.stdout.strip().startswith("ok:"), p.stdout[-500:] + p.stderr[-1500:]


def test_shm_quickstart_example_runs():
    p = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "quickstart_shm.py")], capture_output=True,
                       text=True, timeout=120, env=dict(os.environ, PYTHONPATH=ROOT))
    assert p.returncode == 0 and p.stdout.strip().startswith("ok:"), p.stdout[-500:] + p.stderr[-1500:]


def test_native_cpu_cluster_binary():
    """The host runtime stands alone: coordinator + workers in one native process (no Python in the loop),
    reference defaults (2 workers, dataSize 10, maxChunkSize 2, th 1/1/0.8) and the 4-worker manual config with
    the assertMultiple check."""
    from akka_allreduce_b200.build import build_cpu_cluster

    exe = str(build_cpu_cluster())
    for argv in (["2", "10", "2", "500", "1", "1", "1", "0.8"], ["4", "778", "3", "200", "3", "1", "1", "1"]):
        p = subprocess.run([exe] + argv, capture_output=True, text=True, timeout=120)
        assert p.returncode == 0 and "finished:" in p.stdout and "0 sink assertion failures" in p.stdout, p.stdout + p.stderr


def test_every_config_key_is_read_somewhere():
    """A key that parses but is never consulted is a silent no-op (device.algo was one): every field of every
    section must be referenced by package code outside config.py."""
    import dataclasses
    import re

    from akka_allreduce_b200 import config as c

    pkg = os.path.join(ROOT, "akka_allreduce_b200")
    text = ""
    for base, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py") and f != "config.py":
                text += open(os.path.join(base, f)).read()
    unused = [f"{sec}.{f.name}" for sec, cls in c._SECTIONS.items() for f in dataclasses.fields(cls)
              if not re.search(rf"\.{f.name}\b", text)]
    assert not unused, unused


def test_cli_reads_the_config_file_without_being_told(tmp_path, monkeypatch):
    """The reference picks up application.conf on its own; so do the launchers (AAR_CONFIG names another file)."""
    p = tmp_path / "other.conf"
    p.write_text("allreduce { worker { totalSize = 5, maxLag = 2 }, thresholds.thComplete = 0.5 }")
    monkeypatch.setenv("AAR_CONFIG", str(p))
    c = build_config([])
    assert (c.worker.totalSize, c.worker.maxLag, c.thresholds.thComplete) == (5, 2, 0.5)
    assert build_config(["2600", "3"]).worker.totalSize == 3          # the command line still wins
    monkeypatch.delenv("AAR_CONFIG")
    assert build_config([]).worker.totalSize == 2                     # conf/application.conf


@pytest.mark.timeout(120)
def test_workers_survive_garbage_on_their_ports():
    """Random frames of every kind (valid lengths, random payloads) at the workers' own listening ports while a job
    runs: control frames are only taken from the coordinator's connection, absurd round numbers are dropped, and
    the job keeps making progress and shuts down cleanly.  (Before those two checks a random StartAllreduce frame
    sent the worker into a catch-up loop of billions of rounds.)"""
    import random
    import socket
    import struct

    rng = random.Random(5)
    cfg = _cluster(2, 64, 8, 10 ** 8)
    master = AllreduceMaster(cfg)
    mport = master.start()
    workers = []
    for _ in range(2):
        w = AllreduceWorker(lambda req: AllReduceInput(np.ones(64, dtype=np.float32)), lambda out: None,
                            master_port=mport)
        w.start()
        workers.append(w)
    try:
        t0 = time.time()
        while master.round < 10 and time.time() - t0 < 20:
            time.sleep(0.01)
        ports = [w._node.port for w in workers]
        for _ in range(150):
            try:
                s = socket.create_connection(("127.0.0.1", rng.choice(ports)), timeout=2)
                for _ in range(rng.randint(1, 20)):
                    n = rng.choice([0, 1, 3, 4, 7, 16, 40, 200, 5000])
                    s.sendall(struct.pack("<IB", n, rng.randint(0, 12)) + os.urandom(n))
                s.close()
            except OSError:
                pass
        r0 = master.round
        t0 = time.time()
        while master.round < r0 + 50 and time.time() - t0 < 20:
            time.sleep(0.01)
        assert master.round >= r0 + 50 and master.alive == 2, (master.round, r0, master.alive)
    finally:
        t0 = time.time()
        for w in workers:
            w.stop()
        master.stop()
    assert time.time() - t0 < 10          # stop() does not hang on a wedged IO thread


def test_handler_mode_worker_node_drives_the_coordinator():
    """The plumbing `DeviceWorker` uses, without a GPU: a WorkerNode with a Python message handler instead of an
    engine.  InitWorkers hands out the id, every StartAllreduce is answered with report_complete(id, round), and the
    coordinator (which only counts a report from the connection holding that id) runs through all rounds."""
    import threading

    from akka_allreduce_b200 import host

    cfg = _cluster(3, 16, 4, 40)
    master = AllreduceMaster(cfg)
    port = master.start()
    nodes, state = [], []
    for i in range(3):
        opts = host.WorkerNodeOptions()
        node = host.WorkerNode("127.0.0.1", port, None, None, opts)
        st = {"id": -1, "started": -1, "reported": -1, "lock": threading.Lock()}

        def on_message(msg, node=node, st=st):
            with st["lock"]:
                if isinstance(msg, host.InitWorkers) and st["id"] < 0:
                    st["id"] = msg.dest_id
                    assert msg.worker_num == 3 and msg.data_size == 16 and msg.max_chunk_size == 4
                elif isinstance(msg, host.StartAllreduce):
                    st["started"] = max(st["started"], msg.round)
                if st["id"] >= 0:
                    while st["reported"] < st["started"]:
                        st["reported"] += 1
                        node.report_complete(st["id"], st["reported"])

        node.set_handler(on_message)
        node.listen("127.0.0.1", 0)
        node.begin()
        nodes.append(node)
        state.append(st)
    try:
        assert master.wait(30.0), f"master stuck at round {master.round}"
        assert sorted(s["id"] for s in state) == [0, 1, 2]
        assert all(s["reported"] == 40 for s in state)
        # a report for somebody else's id is not counted (and does not disturb anything)
        nodes[0].report_complete(state[1]["id"], 41)
    finally:
        for n in nodes:
            n.stop()
        master.stop()


def test_reports_survive_a_worker_that_leaves_with_unread_input():
    """Workers that gate rounds themselves run ahead of the coordinator and may leave the moment their last report is
    queued — with StartAllreduce frames they never read still in their socket.  A plain close() then answers with a
    reset, and a reset lets the coordinator's stack drop reports it had received but not yet read (seen as a
    coordinator stuck a few rounds before the end, once in ~15 runs).  stop() half-closes and drains instead."""
    from akka_allreduce_b200 import host

    rounds = 1500
    for attempt in range(4):
        cfg = _cluster(3, 16, 4, rounds)
        master = AllreduceMaster(cfg)
        port = master.start()
        nodes = []
        for _ in range(3):
            node = host.WorkerNode("127.0.0.1", port, None, None, host.WorkerNodeOptions())
            st = {"id": -1, "sent": False}

            def on_message(msg, node=node, st=st):
                if isinstance(msg, host.InitWorkers) and st["id"] < 0:
                    st["id"] = msg.dest_id
                elif isinstance(msg, host.StartAllreduce) and not st["sent"] and st["id"] >= 0:
                    st["sent"] = True
                    for r in range(rounds + 1):          # free-running worker: every report at once ...
                        node.report_complete(st["id"], r)

            node.set_handler(on_message)
            node.listen("127.0.0.1", 0)
            node.begin()
            nodes.append((node, st))
        try:
            t0 = time.time()
            while not all(st["sent"] for _, st in nodes) and time.time() - t0 < 20:
                time.sleep(0.001)
            for node, _ in nodes:                         # ... and gone, without reading what the coordinator sends
                node.stop()
            assert master.wait(30.0), f"attempt {attempt}: coordinator stuck at round {master.round} of {rounds}"
        finally:
            master.stop()


def test_prometheus_metrics_of_master_and_worker():
    """Observability beyond the reference's printed MBytes/sec (its cluster-metrics extension is loaded and never
    read): the coordinator and any worker can be scraped."""
    import urllib.request

    from akka_allreduce_b200.utils.metrics import render
    from akka_allreduce_b200.worker import serve_worker_metrics

    assert render({"aar_x": 3}, {"role": "master"}) == '# TYPE aar_x gauge\naar_x{role="master"} 3\n'
    cfg = _cluster(2, 16, 4, 30)
    master = AllreduceMaster(cfg)
    port = master.start()
    mport = master.serve_metrics(0)
    workers = [AllreduceWorker.builtin(16, checkpoint=10, assert_multiple=2, master_port=port) for _ in range(2)]
    exporters = []
    try:
        for w in workers:
            w.start()
            exporters.append(serve_worker_metrics(w))
        assert master.wait(30)

        def scrape(p):
            text = urllib.request.urlopen(f"http://127.0.0.1:{p}/metrics", timeout=5).read().decode()
            return {l.split("{")[0]: float(l.rsplit(" ", 1)[1]) for l in text.splitlines() if l and not l.startswith("#")}

        m = scrape(mport)
        assert m["aar_master_round"] == 30 and m["aar_master_finished"] == 1 and m["aar_master_total_workers"] == 2
        ws = [scrape(e.port) for e in exporters]
        assert sorted(w["aar_worker_id"] for w in ws) == [0, 1]
        assert all(w["aar_worker_rounds_completed"] == 31 and w["aar_worker_sink_assertion_failures"] == 0 for w in ws)
        assert all(w["aar_worker_checkpoints"] == 3 for w in ws)
        with pytest.raises(Exception):
            urllib.request.urlopen(f"http://127.0.0.1:{mport}/nope", timeout=5)
    finally:
        for e in exporters:
            e.close()
        for w in workers:
            w.stop()
        master.stop()
