"""The MATLAB-subset interpreter (tests/golden/_matlab_subset.py) is what executes the reference's .m text for the golden
vectors and the whole-script run; these cases pin the MATLAB semantics it has to get right (column-major linear indices, `end`,
deletion, repmat, implicit expansion, cells / deal, structs, control flow, user functions with several outputs).  Expected
values are MATLAB's documented behaviour, written out by hand."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import _matlab_subset as ms  # noqa: E402


def run(src, ns=None, functions=None):
    it = ms.Interp(ns if ns is not None else {}, functions or {})
    it.run(src)
    return it.ns


def test_linear_indices_are_column_major_and_end_works():
    ns = run("""
a = [1, 2; 3, 4; 5, 6];
f = a(2);
g = a(4);
h = a(end, 1);
k = a(:, end);
m = a(2: end, :);
n = [a(:, 1: end/2), a(:, end/2+1: end)];
""")
    assert ns["f"] == 3.0 and ns["g"] == 2.0 and ns["h"] == 5.0
    assert np.array_equal(np.ravel(ns["k"]), [2, 4, 6]) and np.array_equal(ns["m"], [[3, 4], [5, 6]])
    assert np.array_equal(ns["n"], ns["a"])


def test_row_deletion_repmat_and_value_semantics():
    ns = run("""
a = [1, 2; 3, 4; 5, 6];
b = a;
b(2, :) = [];
c = repmat(a(1, :), 2, 1);
d = 2.*[10, 20].*c - [1, 1].^2;
""")
    assert np.array_equal(ns["a"], [[1, 2], [3, 4], [5, 6]])            # assignment copies
    assert np.array_equal(ns["b"], [[1, 2], [5, 6]])
    assert np.array_equal(ns["c"], [[1, 2], [1, 2]])
    assert np.array_equal(ns["d"], [[19, 79], [19, 79]])                # 1 x 2 against 2 x 2: implicit expansion


def test_cells_deal_structs_and_control_flow():
    ns = run("""
p.n = 3;
p.v = [4; 8];
c{p.n - 1, 2} = [];
[c{:}] = deal(zeros(1, 2));
c{2, 1} = [7, 9];
s = 0;
for i = 1: p.n
    if i == 2
        continue;
    end
    s = s + i*p.v(end);
    if s > 100
        break;
    end
end
[r, q] = find([0, 1; 1, 0] == 1);
t = sum([1, 2; 3, 4], 'all');
""")
    assert ns["s"] == 8 * (1 + 3)
    assert np.array_equal(ns["c"][(2, 1)], [[7, 9]]) and np.array_equal(ns["c"][(1, 2)], [[0, 0]])
    c2 = ns["c"][(1, 1)]
    c2[0, 0] = 5.0                                                       # deal hands out copies
    assert ns["c"][(1, 2)][0, 0] == 0.0
    assert np.array_equal(np.ravel(ns["r"]), [2, 1]) and np.array_equal(np.ravel(ns["q"]), [1, 2])    # column-major order of find
    assert ns["t"] == 10.0


def test_user_functions_with_several_outputs_and_nargin():
    src = """
function [a, b] = two(x, y)
    if nargin == 1
        y = 10;
    end
    a = x + y;
    b = x*y;
end

function r = outer(x)
    [p, q] = two(x);
    [~, r2] = two(x, 3);
    r = p + q + r2;
end
"""
    fns = ms.parse_functions(src)
    it = ms.Interp({}, fns)
    assert float(it.call_user("outer", [2.0], 1)[0]) == (2 + 10) + 2 * 10 + 2 * 3
    a, b = it.call_user("two", [np.array([[1.0, 2.0]]), 2.0], 2)
    assert np.array_equal(a, [[3, 4]]) and np.array_equal(b, [[2, 4]])
