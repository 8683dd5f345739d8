"""Argument / folder checks used by the power run (mirror of the two helpers `nds_power.py`
imports from the reference's check.py: /root/reference/nds/check.py:136-152, plus check_version
:39-46).  Data-generation helpers of that file are out of scope."""
import os
import sys


def check_version():
    """reference check.py:39-46 requires Python >= 3.6"""
    if sys.version_info < (3, 6):
        raise Exception('Minimum required Python version is 3.6, current is %d.%d.' % sys.version_info[:2])


def check_json_summary_folder(json_summary_folder):
    """Create the folder when missing; refuse a folder that already holds files
    (reference check.py:136-145)."""
    if not json_summary_folder:
        return
    if not os.path.exists(json_summary_folder):
        os.makedirs(json_summary_folder)
    elif os.listdir(json_summary_folder):
        raise Exception(f"json_summary_folder {json_summary_folder} is not empty. " +
                        "There may be already some json files there. Please clean the folder " +
                        "or specify another one.")


def check_query_subset_exists(query_dict, subset_list):
    """Every requested query name must be a key of query_dict (reference check.py:147-152)."""
    for q in subset_list:
        if q not in query_dict.keys():
            raise Exception(f"Query {q} is not in the query dictionary. Please check the query subset.")
    return True
