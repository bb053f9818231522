"""Configuration: one file for both roles, with the reference's key names.

The reference keeps only Akka settings in ``application.conf`` (src/main/resources/
application.conf:1-34) and hard-codes the allreduce parameters in the master launcher
(AllreduceMaster.scala:98-107; README.md:5 promises to move them into the file).  Here they
live in the file.  The typed carriers keep the reference's names — ``ThresholdConfig``,
``DataConfig``, ``WorkerConfig`` (AllreduceMaster.scala:148-150) — and the key names stay
``thAllreduce``, ``thReduce``, ``thComplete``, ``dataSize``, ``maxChunkSize``, ``maxRound``,
``totalSize``, ``maxLag``, ``port``, ``checkpoint``, ``assertMultiple``.

The file format is the HOCON subset the reference's config uses: nested ``{}`` blocks, dotted
keys, ``=`` or ``:``, ``#``/``//`` comments, quoted or bare strings, lists, durations
(``10s``, ``500ms``).  ``.json`` files are read as JSON.
"""
from __future__ import annotations

import json
import os
import re
from dataclasses import dataclass, field, asdict
from pathlib import Path
from typing import Any


# ---------------------------------------------------------------- typed carriers
@dataclass
class ThresholdConfig:
    thAllreduce: float = 1.0   # fraction of workers that must finish round r before r+1 starts
    thReduce: float = 1.0      # fraction of peers whose chunk must arrive before it is reduced
    thComplete: float = 0.8    # fraction of reduced chunks needed to complete a round


@dataclass
class DataConfig:
    dataSize: int = 10         # elements per message
    maxChunkSize: int = 2      # elements per chunk (pipelining / flag granularity)
    maxRound: int = 100        # rounds 0..maxRound inclusive are run


@dataclass
class WorkerConfig:
    totalSize: int = 2         # number of workers the master waits for
    maxLag: int = 1            # rounds a worker may fall behind; ring holds maxLag+1 rounds
    checkpoint: int = 50       # throughput print interval of the built-in sink
    assertMultiple: int = 0    # >0: built-in sink asserts output == input*k, count == k


@dataclass
class ClusterConfig:
    host: str = "127.0.0.1"
    port: int = 2551           # coordinator port (reference default seed node)
    downAfter: float = 10.0    # seconds (auto-down-unreachable-after = 10s)
    announceRemovals: bool = False  # re-send InitWorkers (peer map refresh) when a worker is removed; always on for cuda and shm
    heartbeat: float = 1.0
    loglevel: str = "INFO"
    metricsPort: int = 0       # > 0: the master serves Prometheus metrics on this port (utils/metrics.py)


@dataclass
class DeviceConfig:
    backend: str = "cpu"       # cpu (TCP) | shm (device protocol in shared memory, one host) | cuda
    algo: str = "auto"         # auto | oneshot | twoshot | nvls
    dtype: str = "float32"     # wire dtype of the symmetric slots: float32 | bfloat16
    normalize: bool = False    # divide by the contributor count in the kernel epilogue
    scale: float = 1.0         # source scale fused into the scatter
    gate: str = "device"       # device: thAllreduce gate in symmetric memory; host: via coordinator
    ctas: int = 0              # 0 = pick from message size
    timeoutMs: float = 2000.0  # bound on any in-kernel flag wait


@dataclass
class AllreduceConfig:
    thresholds: ThresholdConfig = field(default_factory=ThresholdConfig)
    data: DataConfig = field(default_factory=DataConfig)
    worker: WorkerConfig = field(default_factory=WorkerConfig)
    cluster: ClusterConfig = field(default_factory=ClusterConfig)
    device: DeviceConfig = field(default_factory=DeviceConfig)

    def to_dict(self) -> dict:
        return asdict(self)

    def master_config(self):
        """The C++ ``MasterConfig`` for this configuration."""
        from . import host

        c = host.MasterConfig()
        c.total_workers = self.worker.totalSize
        c.th_allreduce = self.thresholds.thAllreduce
        c.th_reduce = self.thresholds.thReduce
        c.th_complete = self.thresholds.thComplete
        c.max_lag = self.worker.maxLag
        c.data_size = self.data.dataSize
        c.max_round = self.data.maxRound
        c.max_chunk_size = self.data.maxChunkSize
        return c


# ---------------------------------------------------------------- HOCON subset
_TOKEN = re.compile(
    r"""\s*(?:
        (?P<comment>(?:\#|//)[^\n]*) |
        (?P<brace>[{}\[\],=:]) |
        "(?P<qstr>(?:[^"\\]|\\.)*)" |
        (?P<bare>(?:\$\{[^}\n]*\}|[^\s{}\[\],=:"#$])+)
    )""",
    re.X,
)
_DURATION = re.compile(r"^(-?\d+(?:\.\d+)?)(ms|s|m|h)$")
_UNITS = {"ms": 1e-3, "s": 1.0, "m": 60.0, "h": 3600.0}


_SUBST = re.compile(r"\$\{([^}]*)\}")


def _substitute(tok: str) -> str:
    """``${name}`` as in HOCON: environment variables, plus the JVM properties the reference's file uses."""
    def value(m):
        name = m.group(1).lstrip("?")
        if name == "user.dir":
            return os.getcwd()
        if name == "user.home":
            return os.path.expanduser("~")
        return os.environ.get(name, m.group(0))

    return _SUBST.sub(value, tok)


def _scalar(tok: str) -> Any:
    if "${" in tok:
        tok = _substitute(tok)
    low = tok.lower()
    if low in ("true", "on", "yes"):
        return True
    if low in ("false", "off", "no"):
        return False
    if low == "null":
        return None
    try:
        return int(tok)
    except ValueError:
        pass
    try:
        return float(tok)
    except ValueError:
        pass
    m = _DURATION.match(tok)
    if m:
        return float(m.group(1)) * _UNITS[m.group(2)]
    return tok


def _tokens(text: str):
    pos = 0
    while pos < len(text):
        m = _TOKEN.match(text, pos)
        if not m:
            if text[pos:].strip() == "":
                return
            raise ValueError(f"config: cannot parse near {text[pos:pos + 30]!r}")
        pos = m.end()
        if m.group("comment") is not None:
            continue
        if m.group("brace"):
            yield ("p", m.group("brace"))
        elif m.group("qstr") is not None:
            yield ("s", m.group("qstr"))
        else:
            yield ("b", m.group("bare"))


def _set(tree: dict, dotted: str, value: Any) -> None:
    keys = dotted.split(".")
    for k in keys[:-1]:
        nxt = tree.get(k)
        if not isinstance(nxt, dict):
            nxt = tree[k] = {}
        tree = nxt
    if isinstance(value, dict) and isinstance(tree.get(keys[-1]), dict):
        _merge(tree[keys[-1]], value)
    else:
        tree[keys[-1]] = value


def _merge(dst: dict, src: dict) -> None:
    for k, v in src.items():
        if isinstance(v, dict) and isinstance(dst.get(k), dict):
            _merge(dst[k], v)
        else:
            dst[k] = v


def parse_hocon(text: str) -> dict:
    toks = list(_tokens(text))
    i = 0

    def value():
        nonlocal i
        if i >= len(toks):
            raise ValueError("config: unexpected end of input (a key without a value?)")
        kind, tok = toks[i]
        if (kind, tok) == ("p", "{"):
            i += 1
            return obj(closing=True)
        if (kind, tok) == ("p", "["):
            i += 1
            out = []
            while True:
                if i >= len(toks):
                    raise ValueError("config: missing ']'")
                if toks[i] == ("p", "]"):
                    break
                if toks[i] == ("p", ","):
                    i += 1
                    continue
                out.append(value())
            i += 1
            return out
        if kind == "p":
            raise ValueError(f"config: expected a value, found {tok!r}")
        i += 1
        return tok if kind == "s" else _scalar(tok)

    def obj(closing: bool):
        nonlocal i
        tree: dict = {}
        while i < len(toks):
            kind, tok = toks[i]
            if (kind, tok) == ("p", "}"):
                if not closing:
                    raise ValueError("config: unbalanced '}'")
                i += 1
                return tree
            if (kind, tok) == ("p", ","):
                i += 1
                continue
            if kind == "p":
                raise ValueError(f"config: unexpected {tok!r}")
            key = tok
            i += 1
            if i < len(toks) and toks[i] in (("p", "="), ("p", ":")):
                i += 1
            _set(tree, key, value())
        if closing:
            raise ValueError("config: missing '}'")
        return tree

    return obj(closing=False)


# ---------------------------------------------------------------- load / override
_SECTIONS = {"thresholds": ThresholdConfig, "data": DataConfig, "worker": WorkerConfig,
             "cluster": ClusterConfig, "device": DeviceConfig}
# Flat aliases so a key can be given without its section (and under the reference's names).
_ALIASES = {"numOfWorkers": ("worker", "totalSize"), "totalWorkers": ("worker", "totalSize"),
            "down-after": ("cluster", "downAfter"), "auto-down-unreachable-after": ("cluster", "downAfter")}
for _sec, _cls in _SECTIONS.items():
    for _f in _cls.__dataclass_fields__:
        _ALIASES.setdefault(_f, (_sec, _f))


def _apply(cfg: AllreduceConfig, tree: dict) -> None:
    tree = tree.get("allreduce", tree)
    for key, val in tree.items():
        if key in _SECTIONS and isinstance(val, dict):
            section = getattr(cfg, key)
            for k, v in val.items():
                sec, name = _ALIASES.get(k, (key, k))
                target = getattr(cfg, sec) if sec != key else section
                if name not in target.__dataclass_fields__:
                    raise KeyError(f"config: unknown key {key}.{k}")
                setattr(target, name, type(getattr(target, name))(v))
        elif key in _ALIASES:
            sec, name = _ALIASES[key]
            target = getattr(cfg, sec)
            setattr(target, name, type(getattr(target, name))(val))
        elif key == "akka":
            # Accept the reference's own file: pick up the port, loglevel and auto-down timeout.
            tcp = val.get("remote", {}).get("netty", {}).get("tcp", {}) if isinstance(val, dict) else {}
            if tcp.get("port"):
                cfg.cluster.port = int(tcp["port"])
            if tcp.get("hostname"):
                cfg.cluster.host = str(tcp["hostname"])
            if isinstance(val, dict) and "loglevel" in val:
                cfg.cluster.loglevel = str(val["loglevel"])
            seeds = val.get("cluster", {}).get("seed-nodes") if isinstance(val, dict) else None
            if seeds and not tcp.get("port"):
                # the first seed node is where the reference's master listens (akka.tcp://ClusterSystem@host:2551)
                m = re.search(r"@([^:/]+):(\d+)", str(seeds[0]))
                if m:
                    cfg.cluster.host, cfg.cluster.port = m.group(1), int(m.group(2))
            down = val.get("cluster", {}).get("auto-down-unreachable-after") if isinstance(val, dict) else None
            if down is not None:
                cfg.cluster.downAfter = float(down)
        else:
            raise KeyError(f"config: unknown key {key}")


def load_config(path: str | Path | None = None, overrides: dict | None = None) -> AllreduceConfig:
    """Reads ``path`` (HOCON subset or JSON) and applies ``overrides`` (flat or sectioned keys)."""
    cfg = AllreduceConfig()
    if path is not None:
        text = Path(path).read_text()
        tree = json.loads(text) if str(path).endswith(".json") else parse_hocon(text)
        _apply(cfg, tree)
    if overrides:
        _apply(cfg, {k: v for k, v in overrides.items() if v is not None})
    return cfg


# Two identical copies, like the reference's src/main/resources/application.conf and conf/application.conf:
# the one in conf/ is for editing in a checkout, the packaged one travels with an installed package.
PACKAGED_CONFIG_PATH = Path(__file__).resolve().parent / "resources" / "application.conf"
_checkout_conf = Path(__file__).resolve().parent.parent / "conf" / "application.conf"
DEFAULT_CONFIG_PATH = _checkout_conf if _checkout_conf.exists() else PACKAGED_CONFIG_PATH
