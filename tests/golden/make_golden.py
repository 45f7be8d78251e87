#!/usr/bin/env python3
"""Extract the fixed inputs of the reference's own tests into small fixtures.

  ur5_random_numbers_m1.npy  <- /root/reference/tests/random-numbers.hh  (matrix m1, 220x6:
      the UR5 configurations used by tests/test-continuous-validation.cc:131-176)
  intervals_known_answers.json <- the known-answer cases of /root/reference/tests/test-intervals.cc:55-233
      (sequence of unions -> expected list), parsed from the test source.

Run in the dev container (reads /root/reference, which does not exist on the GPU box).
"""
import json
import os
import re

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def random_numbers():
    src = open(os.path.join(REF, "tests", "random-numbers.hh")).read()
    m = re.search(r"m1\s*<<(.*?);", src, re.S)
    vals = [float(x) for x in re.findall(r"[-+]?\d*\.?\d+(?:[eE][-+]?\d+)?", m.group(1))]
    a = np.array(vals, dtype=np.float64)
    assert a.size == 220 * 6, a.size
    np.save(os.path.join(HERE, "ur5_random_numbers_m1.npy"), a.reshape(220, 6))


def intervals_cases():
    src = open(os.path.join(REF, "tests", "test-intervals.cc")).read()
    cases = []
    for name, body in re.findall(r"BOOST_AUTO_TEST_CASE\((intervals?_\d)\)\s*\{(.*?)\n\}", src, re.S):
        unions = [(float(a), float(b)) for a, b in re.findall(r"interval(?:_t interval| = interval_t)\(([-\d.]+),\s*([-\d.]+)\)", body)]
        expect = [float(x) for x in re.findall(r"BOOST_CHECK_EQUAL\(it->(?:first|second),\s*([-\d.]+)\)", body)]
        size = re.search(r"list\(\)\.size\(\)\s*==\s*(\d+)", body)
        if size and expect:
            cases.append(dict(name=name, unions=unions, expected=[[expect[2 * i], expect[2 * i + 1]] for i in range(int(size.group(1)))]))
    with open(os.path.join(HERE, "intervals_known_answers.json"), "w") as f:
        json.dump(cases, f, indent=1)
    return cases


if __name__ == "__main__":
    random_numbers()
    c = intervals_cases()
    print("wrote fixtures;", len(c), "interval cases")
