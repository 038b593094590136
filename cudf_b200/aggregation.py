"""Aggregation factories with pylibcudf's names (/root/reference/python/pylibcudf/pylibcudf/aggregation.pyx)."""
from __future__ import annotations

from dataclasses import dataclass


class Kind:
    """cudf::aggregation::Kind (/root/reference/cpp/include/cudf/aggregation.hpp:78-121)."""
    SUM, SUM_OVERFLOW, PRODUCT, MIN, MAX, COUNT_VALID, COUNT_ALL, ANY, ALL, SUM_OF_SQUARES = range(10)
    MEAN, M2, VARIANCE, STD, MEDIAN, QUANTILE, ARGMAX, ARGMIN, NUNIQUE, NTH_ELEMENT = range(10, 20)
    MERGE_COUNT = 100  # library extension: sum of INT32 partial counts -> INT32 (include/bgb.h)


class NullPolicy:
    EXCLUDE = 0
    INCLUDE = 1


@dataclass(frozen=True)
class Aggregation:
    kind: int
    ddof: int = 1


def sum(): return Aggregation(Kind.SUM)  # noqa: A001,E704
def sum_overflow(): return Aggregation(Kind.SUM_OVERFLOW)  # noqa: E704  (struct{sum, overflow} result: Column.children)
def product(): return Aggregation(Kind.PRODUCT)  # noqa: E704
def min(): return Aggregation(Kind.MIN)  # noqa: A001,E704
def max(): return Aggregation(Kind.MAX)  # noqa: A001,E704
def sum_of_squares(): return Aggregation(Kind.SUM_OF_SQUARES)  # noqa: E704
def mean(): return Aggregation(Kind.MEAN)  # noqa: E704
def m2(): return Aggregation(Kind.M2)  # noqa: E704
def argmin(): return Aggregation(Kind.ARGMIN)  # noqa: E704
def argmax(): return Aggregation(Kind.ARGMAX)  # noqa: E704
def median(): return Aggregation(Kind.MEDIAN)  # noqa: E704  (not a hash aggregation: UnsupportedError)
def any(): return Aggregation(Kind.ANY)  # noqa: A001,E704
def all(): return Aggregation(Kind.ALL)  # noqa: A001,E704


def count(null_handling: int = NullPolicy.EXCLUDE):
    """make_count_aggregation: COUNT_VALID for EXCLUDE, COUNT_ALL for INCLUDE (cudf/aggregation.hpp)."""
    return Aggregation(Kind.COUNT_VALID if null_handling == NullPolicy.EXCLUDE else Kind.COUNT_ALL)


def variance(ddof: int = 1):
    return Aggregation(Kind.VARIANCE, ddof)


def std(ddof: int = 1):
    return Aggregation(Kind.STD, ddof)


def merge_count():
    """Sum of partial COUNT columns, returned as size_type (the reduce step of a distributed / streaming groupby)."""
    return Aggregation(Kind.MERGE_COUNT)
