"""CPU oracle for the hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package.  See oracle.c for the algorithm restatements and their reference citations.
"""
from .pyoracle import *  # noqa: F401,F403
