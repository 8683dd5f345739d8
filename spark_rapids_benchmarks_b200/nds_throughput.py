"""Throughput run: N power-run processes at once, one query stream each, sharing the GPU(s).

Mirror of the reference's `nds-throughput` shell script (/root/reference/nds/nds-throughput:18-23:
`echo -n "$1" | xargs -d ',' -I '{}' -P$num_streams -r -t "${@:2}"`): the first argument is a comma
separated list of stream numbers, the rest is a command in which every `{}` is replaced by one of
them; all commands run concurrently and the exit code is non-zero if any of them failed (xargs: 123).

    python -m spark_rapids_benchmarks_b200.nds_throughput 1,2,3,4 \
        python -m spark_rapids_benchmarks_b200.nds_power /data/nds './streams/query_{}.sql' 'time_{}.csv'

Every process owns its engine, CUDA context and copy of the encoded tables (the reference starts one
Spark application per stream the same way), so nothing is shared between streams but the GPU itself.

The elapsed time of a throughput test follows the reference's nds_bench.py:138-157
(`get_throughput_time`): latest `Power End Time` minus earliest `Power Start Time` over the streams'
time logs, rounded up to 0.1 s; `get_stream_range` (:127-136) picks the streams of the first or the
second throughput test.
"""
import math
import subprocess
import sys


def stream_numbers(arg):
    return [s for s in arg.split(',') if s != '']


def substitute(cmd, stream):
    return [part.replace('{}', stream) for part in cmd]


def run_streams(streams, cmd, echo=True):
    """Start one process per stream, wait for all of them; returns the list of exit codes."""
    procs = []
    for s in streams:
        c = substitute(cmd, s)
        if echo:
            print(' '.join(c), file=sys.stderr, flush=True)       # xargs -t
        procs.append(subprocess.Popen(c))
    return [p.wait() for p in procs]


def round_up_to_nearest_10_percent(num):
    return math.ceil(num * 10) / 10


def get_start_end_time(report_file):
    start_time = end_time = None
    with open(report_file) as f:
        for line in f:
            if 'Power Start Time' in line:
                start_time = line.split(',')[2].strip()
            if 'Power End Time' in line:
                end_time = line.split(',')[2].strip()
    if start_time and end_time:
        return start_time, end_time
    raise Exception('Start or End time not found in Power Test report file: %s' % report_file)


def get_stream_range(num_streams, first_or_second):
    """num_streams counts the power stream too: 9 -> first test [1,2,3,4], second test [5,6,7,8]."""
    if first_or_second == 1:
        return list(range(1, num_streams // 2 + 1))
    return list(range(num_streams // 2 + 1, num_streams))


def get_throughput_time(report_file_base, num_streams, first_or_second):
    starts, ends = [], []
    for n in get_stream_range(num_streams, first_or_second):
        s, e = get_start_end_time(report_file_base + '_%d.csv' % n)
        starts.append(float(s))
        ends.append(float(e))
    return round_up_to_nearest_10_percent(max(ends) - min(starts))


def main(argv=None):
    argv = sys.argv[1:] if argv is None else argv
    if len(argv) < 2:
        print('usage: nds_throughput <stream,stream,...> <command with {} for the stream number>', file=sys.stderr)
        return 2
    streams = stream_numbers(argv[0])
    print(len(streams))
    print(' '.join(argv[1:]))
    codes = run_streams(streams, argv[1:])
    return 0 if all(c == 0 for c in codes) else 123


if __name__ == '__main__':
    sys.exit(main())
