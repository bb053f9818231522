#!/usr/bin/env bash
# Host runtime under the compiler sanitizers (race detection for the C++ side).
#
#   1. the stand-alone native cluster (coordinator + workers over TCP, IO thread + engine threads) built with
#      ASan+UBSan and with TSan, two configurations each (thresholds 1 and 0.7);
#   2. the shared-memory restatement of the DEVICE protocol (csrc/host/shm_cluster.h), free-running with stragglers;
#   3. the pybind11 module built with ASan / TSan and the CPU test files that drive the engines, the bus and the
#      TCP nodes (kill / rejoin / chaos included) run against it (the runtime is pre-loaded into python).
#
# Usage: scripts/sanitize_host.sh [address|thread ...]      (default: both; a few minutes)
set -u
cd "$(dirname "$0")/.."
SANS=("$@"); [ ${#SANS[@]} -eq 0 ] && SANS=(address thread)
CORE="csrc/host/buffers.cc csrc/host/worker_engine.cc csrc/host/master_engine.cc csrc/host/local_bus.cc csrc/host/node.cc"
TESTS="tests/test_protocol_spec.py tests/test_buffers.py tests/test_master_and_cluster.py tests/test_chaos_cpu.py tests/test_config_cli_tcp.py tests/test_shm_cluster.py tests/test_shm_worker.py"
PATTERN='WARNING: ThreadSanitizer|ERROR: AddressSanitizer|runtime error:|ERROR: LeakSanitizer'
mkdir -p build_san
status=0
for san in "${SANS[@]}"; do
  flags="-fsanitize=$san"; [ "$san" = address ] && flags="-fsanitize=address,undefined"
  exe=build_san/cluster_$san
  g++ -std=c++17 -O1 -g -fno-omit-frame-pointer $flags -Iinclude -Icsrc/host $CORE csrc/host/cpu_cluster_main.cc -o $exe -lpthread || exit 1
  for cfg in "4 778 3 200 3 1 1 1" "3 4096 64 300 2 0.7 0.7 0.7"; do
    log=build_san/native_${san}_$(echo $cfg | tr ' ' '_').log
    timeout 600 $exe $cfg > $log 2>&1; rc=$?
    n=$(grep -c -E "$PATTERN" $log)
    echo "[$san] native cluster ($cfg): exit $rc, $n sanitizer reports, $(grep -o '[0-9.]* rounds/s' $log | tail -1)"
    [ $rc -ne 0 ] || [ "$n" -ne 0 ] && status=1
  done
  # (libstdc++ is pre-loaded too: python does not link it, and the runtime must find __cxa_throw when it starts)
  # the device protocol restated for CPU threads (shm_cluster.h): its data words are plain floats, so a read that is
  # not ordered behind its writer by a flag is a race report here
  shm=build_san/shm_$san
  g++ -std=c++17 -O1 -g -fno-omit-frame-pointer $flags -Iinclude -Icsrc/host csrc/host/shm_cluster.cc csrc/host/shm_cluster_main.cc -o $shm -lpthread || exit 1
  for cfg in "4 778 3 3 1000" "8 4096 64 2 600 0.75 0.75 0.75 0.01 100 200" "3 100 7 0 2000 0.7 0.7 0.7 0.01 100 200" "6 512 64 1 1000 0.7 0.7 0.7 0.02 200 20"; do
    log=build_san/shm_${san}_$(echo $cfg | tr ' ' '_').log
    timeout 600 $shm $cfg > $log 2>&1; rc=$?
    n=$(grep -c -E "$PATTERN" $log)
    echo "[$san] shared-memory device protocol ($cfg): exit $rc, $n sanitizer reports, $(grep -o '[0-9.]* rounds/s' $log | tail -1), $(grep -o '[0-9]* mismatches' $log | tail -1)"
    [ $rc -ne 0 ] || [ "$n" -ne 0 ] && status=1
  done
  lib="$(g++ -print-file-name=lib$([ "$san" = thread ] && echo tsan || echo asan).so) $(g++ -print-file-name=libstdc++.so.6)"
  log=build_san/pytest_$san.log
  rm -f build_san/report_$san.*      # (pytest captures stderr: the runtimes write their reports to files)
  AAR_HOST_SANITIZE=$san LD_PRELOAD="$lib" ASAN_OPTIONS=detect_leaks=0:log_path=build_san/report_$san \
    TSAN_OPTIONS="report_signal_unsafe=0:log_path=build_san/report_$san" \
    timeout 1800 python -m pytest $TESTS -q -p no:cacheprovider > $log 2>&1; rc=$?
  n=$(cat build_san/report_$san.* 2>/dev/null | grep -c -E "$PATTERN")
  echo "[$san] pytest against the instrumented module: exit $rc, $n sanitizer reports, $(tail -1 $log)"
  [ $rc -ne 0 ] || [ "$n" -ne 0 ] && status=1
done
exit $status
