"""Whole-benchmark arithmetic: the report parsers and the composite performance metric of the reference's
`nds_bench.py` (/root/reference/nds/nds_bench.py:59-157, 207-208, 334-357).

The orchestration half of that file (data generation, load test, maintenance phases) drives Spark jobs and
Iceberg/Delta warehouses and stays out of scope; what a run of THIS engine can feed are the
power-run and throughput-run time logs and the transcode load report, all of which these functions parse in
the reference's own formats.  Maintenance times (Tdm1 / Tdm2) have no source here, so a caller supplies them
from a Spark run or reports the metric as incomplete.
"""
from .nds_throughput import (get_start_end_time, get_stream_range, get_throughput_time,  # noqa: F401
                             round_up_to_nearest_10_percent)


def _last_field(path, needle, pick, what):
    found = None
    with open(path) as f:
        for line in f:
            if needle in line:
                found = pick(line)
    if not found:
        raise Exception('%s not found in %s' % (what, path))
    return found


def get_load_end_timestamp(load_report_file):
    """'RNGSEED used: 07291122510' -> '07291122510' (the seed handed to the query-stream generator)."""
    return _last_field(load_report_file, 'RNGSEED used:', lambda l: l.split(':')[1].strip(), 'RNGSEED')


def get_load_time(load_report_file):
    """'Load Test Time: 1234 seconds' -> '1234'."""
    return _last_field(load_report_file, 'Load Test Time', lambda l: l.split(':')[1].split(' ')[1], 'Load Test Time')


def get_power_time(power_report_file):
    """'app,Power Test Time,11838' -> '11838' (milliseconds, as nds_power writes it)."""
    return _last_field(power_report_file, 'Power Test Time', lambda l: l.split(',')[2].strip(), 'Power Test Time')


def get_throughput_stream_nums(num_streams, first_or_second):
    return ','.join(str(x) for x in get_stream_range(num_streams, first_or_second))


def get_perf_metric(scale_factor, num_streams_in_throughput, Tload, Tpower, Ttt1, Ttt2, Tdm1, Tdm2):
    """Primary metric of the benchmark: int(SF * Q / (Tpt * Ttt * Tdm * Tld) ** (1/4)) with all times in hours,
    Q = 99 queries per stream, the power time weighted by the stream count and the load time by 1 % of it
    (reference nds_bench.py:334-357; inputs in seconds)."""
    streams = num_streams_in_throughput
    hours = lambda seconds: seconds / 3600
    t_power = hours(Tpower * streams)
    t_throughput = hours(Ttt1 + Ttt2)
    t_maintenance = hours(Tdm1 + Tdm2)
    t_load = hours(0.01 * streams * Tload)
    return int(int(scale_factor) * (streams * 99) / (t_power * t_throughput * t_maintenance * t_load) ** (1 / 4))


def write_metrics_report(report_path, metrics_map):
    """`key,value` lines (reference nds_bench.py:360-364)."""
    with open(report_path, 'w') as f:
        for key, value in metrics_map.items():
            f.write('%s,%s\n' % (key, value))
