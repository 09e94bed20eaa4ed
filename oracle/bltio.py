"""Readers for the reference's on-disk formats.  TEST INFRASTRUCTURE (part of oracle/).

  read_bilp     -- `l <= a1 ... an <= u` lines            (libblt/parser.cc:72-166)
  read_boost    -- Boost ublas text `[r,c]((..),(..))`     (libblt/parser.cc:43-55; the .lat / .con files)
  read_ntl      -- NTL matrix text `[[a b ..][..]]`        (libblt/ntl_wrapper.cc:191-203; lattice_95.tst)
"""
from __future__ import annotations

import re

import numpy as np


def parse_bilp(text: str):
    """Returns (L, C): L is rows x vars (one row per inequality), C is rows x 2.  Raises ValueError where the
    reference returns 1 (parser.cc:98-103, :142-146)."""
    rows, cons = [], []
    nc = 0
    for ln, line in enumerate(text.splitlines()):
        toks = line.split()
        nt = len(toks)
        if nt < 5 or toks[1] != "<=" or toks[nt - 2] != "<=":
            raise ValueError(f"parse error on input line {ln + 1}")
        vals = [int(toks[0])] + [int(t) for t in toks[2:nt - 2]] + [int(toks[nt - 1])]
        if nc and len(vals) - 2 != nc:
            raise ValueError(f"parse error on input line {ln + 1}")
        nc = len(vals) - 2
        cons.append((vals[0], vals[-1]))
        rows.append(vals[1:-1])
    if not rows or nc == 0:
        raise ValueError("input stream is empty")
    return np.array(rows, dtype=np.int64), np.array(cons, dtype=np.int64)


def read_bilp(path: str):
    with open(path) as f:
        return parse_bilp(f.read())


def parse_boost(text: str) -> np.ndarray:
    mt = re.match(r"\s*\[(\d+),(\d+)\]", text)
    if not mt:
        raise ValueError("boost matrix parser failed")
    r, c = int(mt.group(1)), int(mt.group(2))
    nums = [int(t) for t in re.findall(r"-?\d+", text[mt.end():])]
    if len(nums) != r * c:
        raise ValueError("boost matrix parser failed")
    return np.array(nums, dtype=np.int64).reshape(r, c)


def read_boost(path: str) -> np.ndarray:
    with open(path) as f:
        return parse_boost(f.read())


def read_ntl(path: str):
    """Returns a list of lists of Python ints (entries may exceed int64)."""
    with open(path) as f:
        text = f.read()
    rows = re.findall(r"\[([^\[\]]*)\]", text)
    return [[int(t) for t in r.split()] for r in rows if r.strip()]
