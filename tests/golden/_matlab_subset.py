"""A small interpreter for the subset of MATLAB that the outer-loop statements of the reference's CVX script use
(matlab_old_files/ADMM_CVX_two_veh_intesection_PI_antiwindup.m: parameter block :5-25, collision test :105-121, dual / PI
update with saturation and back-calculation :152-188, residual sums and stop test :190-209), so that those source lines can be
EXECUTED VERBATIM in the build container (no MATLAB / Octave here) instead of being transcribed by hand.

Supported: numeric scalars and 2-D matrices (numpy), cell arrays `c{i,j}`, structs `s.f`, 1-based indexing with ranges and
`:`, matrix literals `[a, b; c, d]` (comma / semicolon separated), operators  + - * / ^ .* ./ .^ '  == ~= < <= > >= && || ~ ,
`if / else / end`, `for v = a:b`, `continue`, `break`, `[r, c] = find(M)`, and the functions sqrt, exp, cos, sin, diag, min,
max, sum(x, 'all'), zeros, ones, length.  `*` is the matrix product unless an operand is a scalar.  Anything else raises.
Test infrastructure only (used by tests/golden/make_reference_vectors.py); never a product path."""
import re

import numpy as np

TOKEN = re.compile(r"""
    (?P<num>\d+\.?\d*(?:[eE][-+]?\d+)?|\.\d+) |
    (?P<id>[A-Za-z_]\w*) |
    (?P<op>\.\^|\.\*|\./|==|~=|<=|>=|&&|\|\||[-+*/^'=<>~(){}\[\],;:.]) |
    (?P<ws>[ \t]+) | (?P<nl>\n)
""", re.X)


class Break(Exception):
    pass


class Continue(Exception):
    pass


def tokenize(src):
    out = []
    for line in src.split("\n"):
        # strip comments (a % that is not inside a string; strings here are only 'all')
        cut, in_str, prev = len(line), False, ""
        for i, ch in enumerate(line):
            if ch == "'" and (in_str or not (prev.isalnum() or prev in ")]}'_.")):
                in_str = not in_str
            elif ch == "%" and not in_str:
                cut = i
                break
            if not ch.isspace():
                prev = ch
        line = line[:cut].rstrip()
        cont = line.endswith("...")
        if cont:
            line = line[:-3]
        pos = 0
        prev_tok = None
        while pos < len(line):
            if line[pos] == "'" and not (prev_tok and (prev_tok[0] in ("num", "id") or prev_tok[1] in (")", "]", "}", "'"))):
                end = line.index("'", pos + 1)
                prev_tok = ("str", line[pos + 1:end]); out.append(prev_tok); pos = end + 1
                continue
            m = TOKEN.match(line, pos)
            if not m:
                raise SyntaxError(f"cannot tokenize {line[pos:]!r}")
            pos = m.end()
            kind = m.lastgroup
            if kind == "ws":
                continue
            prev_tok = (kind, m.group())
            out.append(prev_tok)
        if not cont:
            out.append(("nl", "\n"))
    return out


def is_scalar(x):
    return np.ndim(x) == 0


def mat(x):
    x = np.asarray(x, dtype=np.float64)
    return x.reshape(1, 1) if x.ndim == 0 else (x.reshape(1, -1) if x.ndim == 1 else x)


def simplify(x):
    if isinstance(x, np.ndarray) and x.size == 1 and x.ndim <= 2 and not isinstance(x, Cell):
        return float(x.reshape(-1)[0])
    return x


class Cell(dict):
    pass


class Struct(dict):
    pass


NOOP_COMMANDS = ("tic", "hold", "refreshdata", "drawnow", "clear", "rng", "fprintf", "plot", "figure", "disp")


def parse_functions(src):
    """`function [o1, o2] = name(a1, a2) ... end` definitions at the end of a script file -> {name: (outs, args, body)}."""
    out = {}
    lines = src.split("\n")
    i = 0
    while i < len(lines):
        m = re.match(r"\s*function\s+(?:\[([^\]]*)\]|(\w+))\s*=\s*(\w+)\s*\(([^)]*)\)", lines[i])
        if not m:
            i += 1
            continue
        outs = [o.strip() for o in (m.group(1) or m.group(2)).split(",")]
        args = [a.strip() for a in m.group(4).split(",")]
        depth, j = 1, i + 1
        while j < len(lines):
            code = lines[j].split("%")[0]
            depth += len(re.findall(r"\b(?:if|for|while|function)\b", code))
            depth -= len(re.findall(r"(?<![\w:(,-])\bend\b(?!\s*[-+)\],])", code)) if not re.search(r"[({].*\bend\b", code) else 0
            if depth == 0:
                break
            j += 1
        out[m.group(3)] = (outs, args, "\n".join(lines[i + 1:j]))
        i = j + 1
    return out


class Interp:
    def __init__(self, ns=None, functions=None, hooks=None):
        self.ns = ns if ns is not None else {}
        self.functions = functions or {}
        self.hooks = hooks or {}

    def call_user(self, name, argv, nout):
        outs, args, body = self.functions[name]
        local = {a: self.value_copy(v) for a, v in zip(args, argv)}
        local["nargin"] = float(len(argv))
        Interp(local, self.functions, self.hooks).run(body)
        return [local[o] for o in outs[:max(nout, 1)]]

    # ------------------------------------------------------------------ driver
    def run(self, src):
        self.t = tokenize(src)
        self.i = 0
        self.block(("__eof__",))

    def peek(self):
        return self.t[self.i] if self.i < len(self.t) else ("eof", "")

    def next(self):
        tok = self.peek()
        self.i += 1
        return tok

    def accept(self, val):
        if self.peek()[1] == val and self.peek()[0] != "str":
            self.i += 1
            return True
        return False

    def expect(self, val):
        if not self.accept(val):
            raise SyntaxError(f"expected {val!r}, got {self.peek()!r}")

    def skip_seps(self):
        while self.peek()[0] == "nl" or (self.peek()[0] == "op" and self.peek()[1] in (";", ",")):
            self.i += 1

    # ------------------------------------------------------------------ statements (executed while parsing; loops re-parse)
    def block(self, stops, execute=True):
        """Runs statements until one of the keywords in `stops`; returns the keyword."""
        while True:
            self.skip_seps()
            kind, val = self.peek()
            if kind == "eof":
                return "__eof__"
            if kind == "id" and val in stops:
                self.i += 1
                return val
            self.statement(execute)

    def skip_block(self, stops):
        return self.block(stops, execute=False)

    def statement(self, execute):
        kind, val = self.peek()
        if kind == "id" and val == "if":
            self.i += 1
            cond = self.expr(execute)
            taken = execute and self.truth(cond)
            stop = self.block(("else", "elseif", "end"), taken)
            if stop == "elseif":
                raise SyntaxError("elseif is not supported")
            if stop == "else":
                self.block(("end",), execute and not taken)
            return
        if kind == "id" and val == "for":
            self.i += 1
            var = self.next()[1]
            self.expect("=")
            rng = self.expr(execute)
            body = self.i
            if not execute:
                self.skip_block(("end",))
                return
            vals = [rng] if is_scalar(rng) else list(mat(rng).reshape(-1))
            end_i = None
            broke = False
            for v in vals:
                self.i = body
                self.ns[var] = float(v)
                try:
                    self.block(("end",), True)
                    end_i = self.i
                except Continue:
                    pass
                except Break:
                    broke = True
                    break
            self.i = body
            self.skip_block(("end",))
            return
        if kind == "id" and val == "continue":
            self.i += 1
            if execute:
                raise Continue()
            return
        if kind == "id" and val == "break":
            self.i += 1
            if execute:
                raise Break()
            return
        if kind == "id" and val in self.hooks:     # harness hooks standing in for text that cannot run (cvx_begin ... cvx_end)
            self.i += 1
            if execute:
                self.hooks[val](self.ns)
            return
        if kind == "id" and val in NOOP_COMMANDS and val not in self.ns:      # `tic`, `hold on`, `fprintf(...)`, `plot(...)`: to end of line
            while self.peek()[0] not in ("nl", "eof"):
                self.i += 1
            return
        if kind == "op" and val == "[":
            save = self.i
            got = self.try_general_multi(execute)
            if got:
                return
            self.i = save
        if kind == "op" and val == "[":          # [a, b] = find(...)  or an expression statement
            save = self.i
            names = self.try_multi_lhs()
            if names is not None:
                self.expect("=")
                fn = self.next()[1]
                self.expect("(")
                arg = self.expr(execute)
                self.expect(")")
                if execute:
                    if fn != "find":
                        raise NotImplementedError(fn)
                    r, c = np.nonzero(mat(arg).T)          # column-major order, as MATLAB
                    self.ns[names[0]] = simplify(mat((c + 1).astype(float)).T) if len(c) != 1 else float(c[0] + 1)
                    self.ns[names[1]] = simplify(mat((r + 1).astype(float)).T) if len(r) != 1 else float(r[0] + 1)
                return
            self.i = save
        # assignment or expression statement
        save = self.i
        lhs = self.try_lhs(execute)
        if lhs is not None and self.peek() == ("op", "="):
            self.i += 1
            val_ = self.expr(execute)
            if execute:
                self.assign(lhs, val_)
            return
        self.i = save
        self.expr(execute)

    def try_general_multi(self, execute):
        """`[lhs1, lhs2, ~] = f(args)` with indexed targets, a user function or find; `[c{:}] = deal(v)`."""
        start = self.i
        depth, j = 0, self.i
        while True:                                  # find the matching ']' and require '=' after it (not '==')
            k, v = self.t[j]
            if k in ("nl", "eof"):
                return False
            if k == "op" and v in "[({":
                depth += 1
            elif k == "op" and v in "])}":
                depth -= 1
                if depth == 0:
                    break
            j += 1
        if self.t[j + 1] != ("op", "="):
            return False
        self.i += 1
        targets = []
        while True:
            if self.accept("~"):
                targets.append(None)
            else:
                kind, val = self.peek()
                if kind == "id" and self.t[self.i + 1] == ("op", "{") and self.t[self.i + 2] == ("op", ":"):
                    self.i += 4                        # name { : }
                    targets.append(("dealall", val))
                else:
                    targets.append(self.try_lhs(execute))
            if self.accept("]"):
                break
            self.expect(",")
        self.expect("=")
        fn = self.next()[1]
        self.expect("(")
        argv = self.args(")", execute)
        if not execute:
            return True
        if fn == "deal":
            for t in targets:
                cell = self.ns[t[1]]
                n = max(k[0] for k in cell), max(k[1] for k in cell)
                for a in range(1, n[0] + 1):
                    for b in range(1, n[1] + 1):
                        cell[(a, b)] = self.value_copy(argv[0])
            return True
        if fn == "find":
            r, c = np.nonzero(mat(argv[0]).T)
            vals = [simplify(mat((c + 1).astype(float)).T), simplify(mat((r + 1).astype(float)).T)]
        elif fn in self.functions:
            vals = self.call_user(fn, argv, len(targets))
        else:
            raise NotImplementedError(f"multi-output call of {fn}")
        for t, v in zip(targets, vals):
            if t is not None:
                self.assign(t, v)
        return True

    def try_multi_lhs(self):
        self.expect("[")
        names = []
        while True:
            kind, val = self.next()
            if kind != "id":
                return None
            names.append(val)
            if self.accept("]"):
                break
            if not self.accept(","):
                return None
        return names if self.peek() == ("op", "=") else None

    def try_lhs(self, execute):
        kind, val = self.peek()
        if kind != "id":
            return None
        self.i += 1
        path = [("var", val)]
        while True:
            if self.peek() == ("op", "."):
                self.i += 1
                path.append(("field", self.next()[1]))
            elif self.peek() == ("op", "("):
                self.i += 1
                base = self.ns.get(val) if (execute and len(path) == 1) else None
                path.append(("paren", self.args(")", execute, base=base)))
            elif self.peek() == ("op", "{"):
                self.i += 1
                path.append(("brace", self.args("}", execute)))
            else:
                break
        return path

    @staticmethod
    def value_copy(v):
        """MATLAB assigns by VALUE (cell arrays and structs included: `last_iter_hat_pos = hat_pos_old` is a copy)."""
        if isinstance(v, Cell):
            return Cell({k: Interp.value_copy(x) for k, x in v.items()})
        if isinstance(v, Struct):
            return Struct({k: Interp.value_copy(x) for k, x in v.items()})
        return v.copy() if isinstance(v, np.ndarray) else v

    def assign(self, path, value):
        name = path[0][1]
        value = self.value_copy(value)
        if len(path) == 1:
            self.ns[name] = value
            return
        kind, arg = path[1]
        if len(path) != 2:
            raise NotImplementedError("nested assignment")
        if kind == "field":
            self.ns.setdefault(name, Struct())[arg] = value
        elif kind == "brace":
            self.ns.setdefault(name, Cell())[tuple(int(a) for a in arg)] = value
        else:
            base = mat(self.ns[name]).copy() if name in self.ns else np.zeros((0, 0))
            if len(arg) == 1:                                  # linear index into a vector: v(k) = scalar
                k = int(round(arg[0])) - 1
                flat = base.reshape(-1)
                if k >= flat.size:
                    flat = np.concatenate((flat, np.zeros(k + 1 - flat.size)))
                    base = flat.reshape(1, -1) if base.shape[0] <= 1 else flat.reshape(-1, 1)
                base.reshape(-1)[k] = float(value)
                self.ns[name] = base
                return
            if isinstance(value, np.ndarray) and value.size == 0:          # x(i, :) = []  /  x(:, j) = []: deletion
                if isinstance(arg[1], str):
                    self.ns[name] = np.delete(base, int(round(arg[0])) - 1, axis=0)
                else:
                    self.ns[name] = np.delete(base, int(round(arg[1])) - 1, axis=1)
                return
            idx = self.index_tuple(base, arg)
            base[idx] = value if is_scalar(value) else mat(value).reshape(base[idx].shape)
            self.ns[name] = base

    # ------------------------------------------------------------------ expressions
    def truth(self, v):
        return bool(np.all(np.asarray(v) != 0))

    def expr(self, ex):
        return self.or_(ex)

    def or_(self, ex):
        a = self.and_(ex)
        while self.accept("||"):
            b = self.and_(ex)
            a = float(self.truth(a) or self.truth(b)) if ex else 0.0
        return a

    def and_(self, ex):
        a = self.cmp(ex)
        while self.accept("&&"):
            b = self.cmp(ex)
            a = float(self.truth(a) and self.truth(b)) if ex else 0.0
        return a

    def cmp(self, ex):
        a = self.range_(ex)
        while self.peek()[0] == "op" and self.peek()[1] in ("==", "~=", "<", "<=", ">", ">="):
            op = self.next()[1]
            b = self.range_(ex)
            if ex:
                f = {"==": np.equal, "~=": np.not_equal, "<": np.less, "<=": np.less_equal, ">": np.greater, ">=": np.greater_equal}[op]
                a = simplify(f(np.asarray(a, dtype=float), np.asarray(b, dtype=float)).astype(float))
        return a

    def range_(self, ex):
        a = self.add(ex)
        if self.peek() == ("op", ":") and self.range_ok():
            self.i += 1
            b = self.add(ex)
            step = 1.0
            if self.peek() == ("op", ":") and self.range_ok():
                self.i += 1
                step, b = b, self.add(ex)
            if ex:
                return mat(np.arange(int(round(a)), int(round(b)) + (1 if step > 0 else -1), int(round(step)), dtype=float))
            return 0.0
        return a

    def range_ok(self):
        nxt = self.t[self.i + 1] if self.i + 1 < len(self.t) else ("eof", "")
        return not (nxt[0] == "op" and nxt[1] in (",", ")", "}"))        # a bare ':' inside an index list is handled by args()

    def add(self, ex):
        a = self.mul(ex)
        while self.peek()[0] == "op" and self.peek()[1] in ("+", "-"):
            op = self.next()[1]
            b = self.mul(ex)
            if ex:
                a = simplify(np.add(a, b) if op == "+" else np.subtract(a, b))
        return a

    def mul(self, ex):
        a = self.unary(ex)
        while self.peek()[0] == "op" and self.peek()[1] in ("*", "/", ".*", "./"):
            op = self.next()[1]
            b = self.unary(ex)
            if not ex:
                continue
            if op == "*":
                a = simplify(np.multiply(a, b) if is_scalar(a) or is_scalar(b) else mat(a) @ mat(b))
            elif op == "/":
                if not is_scalar(b):
                    raise NotImplementedError("matrix right division")
                a = simplify(np.divide(a, b))
            elif op == ".*":
                a = simplify(np.multiply(a, b))
            else:
                a = simplify(np.divide(a, b))
        return a

    def unary(self, ex):
        if self.accept("-"):
            v = self.unary(ex)
            return simplify(np.negative(v)) if ex else 0.0
        if self.accept("+"):
            return self.unary(ex)
        if self.accept("~"):
            v = self.unary(ex)
            return simplify((np.asarray(v) == 0).astype(float)) if ex else 0.0
        return self.power(ex)

    def power(self, ex):
        a = self.postfix(ex)
        while self.peek()[0] == "op" and self.peek()[1] in ("^", ".^"):
            op = self.next()[1]
            b = self.unary_for_power(ex)
            if ex:
                if op == "^" and not (is_scalar(a) and is_scalar(b)):
                    raise NotImplementedError("matrix power")
                a = simplify(np.power(a, b))
        return a

    def unary_for_power(self, ex):
        if self.accept("-"):
            v = self.postfix(ex)
            return simplify(np.negative(v)) if ex else 0.0
        return self.postfix(ex)

    def arity(self, close):
        """Number of arguments of the index list that starts at the current token (look-ahead, for `end`)."""
        depth, n, i = 0, 1, self.i
        while True:
            k, v = self.t[i]
            if k == "op" and v in "([{":
                depth += 1
            elif k == "op" and v in ")]}":
                if depth == 0:
                    return n
                depth -= 1
            elif k == "op" and v == "," and depth == 0:
                n += 1
            i += 1

    def args(self, close, ex, base=None):
        out = []
        if self.accept(close):
            return out
        self.end_ctx = getattr(self, "end_ctx", [])
        n_args = self.arity(close)
        while True:
            self.end_ctx.append((base, len(out), n_args))
            try:
                done = self._one_arg(out, close, ex)
            finally:
                self.end_ctx.pop()
            if done:
                return out

    def _one_arg(self, out, close, ex):
        if True:
            if self.peek() == ("op", ":") and self.t[self.i + 1][1] in (",", close):
                self.i += 1
                out.append(":")
            elif self.peek()[0] == "str":
                out.append(self.next()[1])
            else:
                out.append(self.expr(ex))
            if self.accept(close):
                return True
            self.expect(",")
            return False

    def index_tuple(self, base, arg):
        def one(a, n):
            if isinstance(a, str) and a == ":":
                return slice(None)
            if is_scalar(a):
                return int(round(a)) - 1
            return (mat(a).reshape(-1).astype(int) - 1)
        if len(arg) == 1:
            raise NotImplementedError("linear indexed assignment")
        return (one(arg[0], base.shape[0]), one(arg[1], base.shape[1]))

    def index(self, base, arg):
        b = mat(base)
        if len(arg) == 1:
            a = arg[0]
            flat = b.reshape(-1, order="F")
            if isinstance(a, str):
                return mat(flat).T
            if is_scalar(a):
                return float(flat[int(round(a)) - 1])
            return simplify(mat(flat[mat(a).reshape(-1).astype(int) - 1]))
        r, c = self.index_tuple(b, arg)
        out = b[np.ix_(np.atleast_1d(np.arange(b.shape[0])[r]), np.atleast_1d(np.arange(b.shape[1])[c]))]
        return simplify(out)

    FUNCS = {"sqrt": np.sqrt, "exp": np.exp, "cos": np.cos, "sin": np.sin}

    def call(self, name, arg):
        if name in self.FUNCS:
            return simplify(self.FUNCS[name](np.asarray(arg[0], dtype=float)))
        if name == "diag":
            m = mat(arg[0])
            return simplify(np.diag(m).reshape(-1, 1) if min(m.shape) > 1 else np.diag(m.reshape(-1)))
        if name in ("min", "max"):
            f2, f1 = (np.minimum, np.min) if name == "min" else (np.maximum, np.max)
            if len(arg) == 2:
                return simplify(f2(np.asarray(arg[0], dtype=float), np.asarray(arg[1], dtype=float)))
            m = mat(arg[0])
            return float(f1(m)) if 1 in m.shape else simplify(f1(m, axis=0, keepdims=True))
        if name == "sum":
            if len(arg) == 2 and arg[1] == "all":
                return float(np.sum(arg[0]))
            m = mat(arg[0])
            return float(m.sum()) if 1 in m.shape else simplify(m.sum(axis=0, keepdims=True))
        if name in ("zeros", "ones"):
            shape = (int(arg[0]), int(arg[1] if len(arg) > 1 else arg[0]))
            return (np.zeros if name == "zeros" else np.ones)(shape)
        if name == "length":
            return float(max(mat(arg[0]).shape)) if np.size(arg[0]) else 0.0
        if name == "repmat":
            return np.tile(mat(arg[0]), (int(arg[1]), int(arg[2])))
        if name == "linspace":
            return mat(np.linspace(arg[0], arg[1], int(round(arg[2]))))
        if name == "size":
            return mat(np.array(mat(arg[0]).shape, dtype=float))
        if name in self.functions:
            return self.call_user(name, arg, 1)[0]
        raise NotImplementedError(f"function {name}")

    def postfix(self, ex):
        kind, val = self.next()
        if kind == "num":
            v = float(val)
        elif kind == "str":
            v = val
        elif kind == "op" and val == "(":
            v = self.expr(ex)
            self.expect(")")
        elif kind == "op" and val == "[":
            v = self.matrix(ex)
        elif kind == "id":
            if val == "pi":
                v = float(np.pi)
            elif val == "toc" and "toc" not in self.ns:
                v = 0.0
            elif val == "end" and getattr(self, "end_ctx", None):
                base, pos, n_args = self.end_ctx[-1]
                b = mat(base) if base is not None and ex else np.zeros((1, 1))
                v = float(b.size if n_args == 1 else b.shape[pos])
            elif self.peek() == ("op", "(") and val not in self.ns:
                self.i += 1
                arg = self.args(")", ex)
                v = self.call(val, arg) if ex else 0.0
            else:
                v = self.ns[val] if ex else 0.0
        else:
            raise SyntaxError(f"unexpected {val!r}")
        while True:
            if self.peek() == ("op", "(") and kind != "num":
                self.i += 1
                arg = self.args(")", ex, base=v if ex else None)
                v = self.index(v, arg) if ex else 0.0
            elif self.peek() == ("op", "{"):
                self.i += 1
                arg = self.args("}", ex)
                v = v[tuple(int(a) for a in arg)] if ex else 0.0
            elif self.peek() == ("op", "."):
                self.i += 1
                f = self.next()[1]
                v = v[f] if ex else 0.0
            elif self.peek() == ("op", "'"):
                self.i += 1
                v = simplify(mat(v).T) if ex else 0.0
            else:
                return v

    def matrix(self, ex):
        rows, cur = [], []
        if self.accept("]"):
            return np.zeros((0, 0))
        while True:
            while self.peek()[0] == "nl":
                self.i += 1
            cur.append(self.expr(ex))
            if self.accept(","):
                continue
            if self.accept(";") or self.peek()[0] == "nl":
                rows.append(cur); cur = []
                while self.peek()[0] == "nl":
                    self.i += 1
                if self.accept("]"):
                    break
                continue
            self.expect("]")
            rows.append(cur)
            break
        if not ex:
            return 0.0
        blocks = []
        for r in rows:
            parts = [mat(x) for x in r if np.size(x)]
            if parts:
                blocks.append(np.hstack(parts))
        return simplify(np.vstack(blocks)) if blocks else np.zeros((0, 0))
