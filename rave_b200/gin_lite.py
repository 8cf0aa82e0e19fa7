"""A small reader for the subset of gin-config syntax the RAVE configs use.

gin-config is not installable here and the reference builds every module through it
(scripts/export.py:547-551 `gin.parse_config_file` then `rave.RAVE()`), so the drop-in
needs to read the same files: macros (`NAME = value`, `%NAME`), `include "file"`,
bindings `scope/module.Class.param = value`, block form (`module.Class:` + indented
params), references `@module.Class`, scoped `@scope/module.Class` and evaluated
`@module.fn()`, python literals (lists, dicts, tuples, numbers, strings, booleans).
`import` lines and `from __gin__ import dynamic_registration` are accepted and ignored.

Selectors are matched on their last identifier (`blocks.EncoderV2` and
`rave.blocks.EncoderV2` both mean `EncoderV2`), which is unambiguous for these files.
"""
import ast
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_SEARCH_PATHS = [os.path.join(_HERE, "configs")]


class Ref:
    """`@scope/module.Name` (evaluate=False) or `@module.name()` (evaluate=True)."""

    def __init__(self, selector, evaluate=False):
        scope, _, name = selector.rpartition("/")
        self.scope = scope
        self.selector = name
        self.name = name.split(".")[-1]
        self.evaluate = evaluate

    def __repr__(self):
        return f"@{self.scope + '/' if self.scope else ''}{self.selector}{'()' if self.evaluate else ''}"


class Macro:
    def __init__(self, name):
        self.name = name

    def __repr__(self):
        return "%" + self.name


_REF_RE = re.compile(r"@([A-Za-z_][\w./]*)(\(\))?")
_MACRO_RE = re.compile(r"%([A-Za-z_]\w*)")


def _strip_comment(line):
    out, quote = [], None
    for ch in line:
        if quote:
            out.append(ch)
            if ch == quote:
                quote = None
        elif ch in "\"'":
            quote = ch
            out.append(ch)
        elif ch == "#":
            break
        else:
            out.append(ch)
    return "".join(out).rstrip()


def _balanced(text):
    depth, quote = 0, None
    for ch in text:
        if quote:
            if ch == quote:
                quote = None
        elif ch in "\"'":
            quote = ch
        elif ch in "([{":
            depth += 1
        elif ch in ")]}":
            depth -= 1
    return depth <= 0


def _literal(node):
    """Value of a literal-only expression tree.  gin itself parses literals only; a config.gin comes from a
    run directory that may have been downloaded, so nothing here is ever evaluated as code: attribute
    access, subscripts, operators other than a numeric sign, and calls other than the two placeholders
    that stand for `@ref` / `%macro` are rejected."""
    if isinstance(node, ast.Constant):
        return node.value
    if isinstance(node, ast.List):
        return [_literal(e) for e in node.elts]
    if isinstance(node, ast.Tuple):
        return tuple(_literal(e) for e in node.elts)
    if isinstance(node, ast.Set):
        return {_literal(e) for e in node.elts}
    if isinstance(node, ast.Dict):
        if any(k is None for k in node.keys):
            raise ValueError("dict unpacking")
        return {_literal(k): _literal(v) for k, v in zip(node.keys, node.values)}
    if isinstance(node, ast.UnaryOp) and isinstance(node.op, (ast.USub, ast.UAdd)):
        v = _literal(node.operand)
        if isinstance(v, bool) or not isinstance(v, (int, float)):
            raise ValueError("sign on a non-number")
        return -v if isinstance(node.op, ast.USub) else v
    if isinstance(node, ast.Call) and isinstance(node.func, ast.Name) and not node.keywords:
        args = [a.value for a in node.args if isinstance(a, ast.Constant)]
        if len(args) == len(node.args):
            if node.func.id == "__REF__" and len(args) == 2:
                return Ref(str(args[0]), bool(args[1]))
            if node.func.id == "__MACRO__" and len(args) == 1:
                return Macro(str(args[0]))
    raise ValueError(f"{type(node).__name__} is not a literal")


def parse_value(text):
    text = text.strip()
    src = _REF_RE.sub(lambda m: f"__REF__({m.group(1)!r}, {bool(m.group(2))})", text)
    src = _MACRO_RE.sub(lambda m: f"__MACRO__({m.group(1)!r})", src)
    try:
        return _literal(ast.parse(src, mode="eval").body)
    except Exception as e:
        raise ValueError(f"gin_lite: cannot parse value {text!r}: {e}")


class Config:
    """Parsed bindings: macros {NAME: value}; bindings {(scope, Name): {param: value}}."""

    def __init__(self, search_paths=None):
        self.macros = {}
        self.bindings = {}
        self.search_paths = list(search_paths or []) + DEFAULT_SEARCH_PATHS
        self.files = []

    # ------------------------------------------------------------------ parsing
    def _resolve(self, path, relative_to=None):
        cands = [path]
        if relative_to:
            cands.append(os.path.join(relative_to, path))
        base = os.path.basename(path)
        for sp in self.search_paths:
            cands += [os.path.join(sp, path), os.path.join(sp, base)]
        if not path.endswith(".gin"):
            cands += [c + ".gin" for c in list(cands)]
        for c in cands:
            if os.path.isfile(c):
                return c
        raise ValueError(f"gin_lite: config file {path!r} not found (searched {self.search_paths})")

    def parse_file(self, path, relative_to=None):
        full = self._resolve(path, relative_to)
        self.files.append(full)
        with open(full) as f:
            self.parse_string(f.read(), os.path.dirname(full))
        return self

    def _bind(self, selector, param, value):
        scope, _, name = selector.rpartition("/")
        key = (scope, name.split(".")[-1])
        self.bindings.setdefault(key, {})[param] = value

    def _assign(self, lhs, rhs, block):
        value = parse_value(rhs)
        lhs = lhs.strip()
        if block is not None:
            self._bind(block, lhs, value)
        elif "." in lhs or "/" in lhs:
            selector, _, param = lhs.rpartition(".")
            self._bind(selector, param, value)
        else:
            self.macros[lhs] = value

    def parse_string(self, text, relative_to=None):
        # gin.operative_config_str() (what train_rave.py:287-288 writes to run/config.gin) wraps long
        # bindings with a trailing backslash: `blocks.EncoderV2.dilations = \` + the value on the next line
        lines, cont = [], None
        for raw in text.split("\n"):
            raw = _strip_comment(raw)
            if cont is not None:
                raw = cont + " " + raw.strip()
                cont = None
            if raw.endswith("\\"):
                cont = raw[:-1].rstrip()
                continue
            lines.append(raw)
        if cont is not None:
            lines.append(cont)
        i, block = 0, None
        while i < len(lines):
            raw = _strip_comment(lines[i])
            i += 1
            if not raw.strip():
                continue
            indented = raw[0] in " \t"
            line = raw.strip()
            if not indented:
                block = None
            if line.startswith(("import ", "from ")):
                continue
            if line.startswith("include"):
                m = re.match(r"include\s+['\"](.+)['\"]", line)
                if not m:
                    raise ValueError(f"gin_lite: bad include line {line!r}")
                self.parse_file(m.group(1), relative_to)
                continue
            if line.endswith(":") and "=" not in line:
                block = line[:-1].strip()
                continue
            if "=" not in line:
                raise ValueError(f"gin_lite: cannot parse line {line!r}")
            lhs, _, rhs = line.partition("=")
            if not rhs.strip() and i < len(lines) and lines[i][:1] in (" ", "\t") and "=" not in lines[i]:
                rhs = lines[i].strip()                      # value alone on the following indented line
                i += 1
            while not _balanced(rhs) and i < len(lines):
                rhs += " " + _strip_comment(lines[i]).strip()
                i += 1
            self._assign(lhs, rhs, block if indented else None)
        return self

    def parse_bindings(self, bindings):
        for b in bindings or []:
            self.parse_string(b)
        return self

    # ------------------------------------------------------------------ queries
    def resolve(self, value):
        """Substitute macros (recursively) inside a parsed value; Refs are kept."""
        if isinstance(value, Macro):
            if value.name not in self.macros:
                raise ValueError(f"gin_lite: macro %{value.name} is not defined")
            return self.resolve(self.macros[value.name])
        if isinstance(value, list):
            return [self.resolve(v) for v in value]
        if isinstance(value, tuple):
            return tuple(self.resolve(v) for v in value)
        if isinstance(value, dict):
            return {k: self.resolve(v) for k, v in value.items()}
        return value

    def params(self, name, scope=""):
        """Bindings for configurable `name`; scoped ones override unscoped ones."""
        name = name.split(".")[-1]
        out = dict(self.bindings.get(("", name), {}))
        if scope:
            parts = scope.split("/")
            for n in range(1, len(parts) + 1):
                out.update(self.bindings.get(("/".join(parts[:n]), name), {}))
                out.update(self.bindings.get((parts[n - 1], name), {}))
        return {k: self.resolve(v) for k, v in out.items()}

    def macro(self, name, default=None):
        if name in self.macros:
            return self.resolve(self.macros[name])
        return default


def parse_config_files_and_bindings(config_files, bindings=None, search_paths=None):
    """Same call shape as gin.parse_config_files_and_bindings (tests/test_configs.py:54-60)."""
    cfg = Config(search_paths)
    if isinstance(config_files, str):
        config_files = [config_files]
    for f in config_files or []:
        cfg.parse_file(f)
    cfg.parse_bindings(bindings)
    return cfg
