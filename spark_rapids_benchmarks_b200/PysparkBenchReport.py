"""Per-query JSON summary - drop-in for the reference's PysparkBenchReport
(/root/reference/nds/PysparkBenchReport.py:42-127): same summary keys, same status strings,
same file name pattern `{prefix}-{query}-{startTime}.json` with indent=2.

Differences forced by the engine swap: `sparkConf` holds the engine session's configuration
dict and `sparkVersion` the engine version string; the py4j task-failure listener
(reference :77-85) has no counterpart (there are no Spark tasks), so the status is either
'Completed' or 'Failed'.
"""
import json
import os
import time
import traceback
from typing import Callable


class PysparkBenchReport:
    """Class to generate json summary report for a benchmark"""

    def __init__(self, spark_session, query_name) -> None:
        self.spark_session = spark_session
        self.summary = {
            'env': {
                'envVars': {},
                'sparkConf': {},
                'sparkVersion': None
            },
            'queryStatus': [],
            'exceptions': [],
            'startTime': None,
            'queryTimes': [],
            'query': query_name,
        }

    def report_on(self, fn: Callable, *args):
        """Run fn(*args), recording environment, wall-clock milliseconds and outcome.
        Environment variables named exactly TOKEN / SECRET / PASSWORD are dropped (reference :70-76).
        The timed region is exactly the call (reference :87-89)."""
        hidden = ("TOKEN", "SECRET", "PASSWORD")
        self.summary['env']['envVars'] = {k: v for k, v in os.environ.items() if k not in hidden}
        self.summary['env']['sparkConf'] = dict(self.spark_session.conf_items())
        self.summary['env']['sparkVersion'] = self.spark_session.version
        start_time = int(time.time() * 1000)
        try:
            fn(*args)
            end_time = int(time.time() * 1000)
            self.summary['queryStatus'].append("Completed")
        except Exception as e:
            # print the exception to ease debugging (reference :94-99)
            print('ERROR BEGIN')
            print(e)
            traceback.print_tb(e.__traceback__)
            print('ERROR END')
            end_time = int(time.time() * 1000)
            self.summary['queryStatus'].append("Failed")
            self.summary['exceptions'].append(str(e))
        self.summary['startTime'] = start_time
        self.summary['queryTimes'].append(end_time - start_time)
        return self.summary

    def write_summary(self, prefix=""):
        """Write the summary as `<prefix>-<query>-<startTime>.json` (reference :110-122)."""
        filename = prefix + '-' + self.summary['query'] + '-' + str(self.summary['startTime']) + '.json'
        self.summary['filename'] = filename
        with open(filename, "w") as f:
            json.dump(self.summary, f, indent=2)

    def is_success(self):
        """queryStatus == Completed (reference :124-127)"""
        return self.summary['queryStatus'][0] == 'Completed'
