"""Flag registry usable both from the command line and as a library.

Same idea as the reference's ``flags.py`` (T/flags.py:36-89): parameters are declared
once as ``ParamSpec`` entries so ``make_params(**kwargs)`` works without a command line,
and ``parse_argv`` turns ``--name=value`` / ``--name value`` / ``--[no]bool`` arguments
into the same structure. There is no absl here; the parser implements the subset of
absl's syntax the reference's scripts and tests use, and rejects stray positional
arguments so that ``--distortions False`` cannot silently mean True
(T/tf_cnn_benchmarks.py:39-46).
"""
from __future__ import annotations

from collections import OrderedDict, namedtuple
from typing import Any, Dict, Sequence

ParamSpec = namedtuple("_ParamSpec", ["flag_type", "default_value", "description", "kwargs"])

# name -> ParamSpec
param_specs: "OrderedDict[str, ParamSpec]" = OrderedDict()
# alias -> canonical name
aliases: Dict[str, str] = {}


def _define(flag_type, name, default, help, **kwargs):
    if name in param_specs:
        raise ValueError("flag %s defined twice" % name)
    param_specs[name] = ParamSpec(flag_type, default, help, kwargs)


def DEFINE_string(name, default, help):
    _define("string", name, default, help)


def DEFINE_boolean(name, default, help):
    _define("boolean", name, default, help)


def DEFINE_integer(name, default, help, lower_bound=None, upper_bound=None):
    kw = {}
    if lower_bound is not None:
        kw["lower_bound"] = lower_bound
    if upper_bound is not None:
        kw["upper_bound"] = upper_bound
    _define("integer", name, default, help, **kw)


def DEFINE_float(name, default, help, lower_bound=None, upper_bound=None):
    kw = {}
    if lower_bound is not None:
        kw["lower_bound"] = lower_bound
    if upper_bound is not None:
        kw["upper_bound"] = upper_bound
    _define("float", name, default, help, **kw)


def DEFINE_enum(name, default, enum_values, help):
    _define("enum", name, default, help, enum_values=list(enum_values))


def DEFINE_list(name, default, help):
    _define("list", name, default, help)


def DEFINE_alias(alias, name):
    aliases[alias] = name


class FlagError(ValueError):
    pass


def _to_bool(text: str) -> bool:
    t = text.strip().lower()
    if t in ("1", "true", "t", "yes", "y"):
        return True
    if t in ("0", "false", "f", "no", "n"):
        return False
    raise FlagError("invalid boolean value %r" % text)


def convert(name: str, text: str) -> Any:
    spec = param_specs[name]
    ft = spec.flag_type
    try:
        if ft == "string":
            return text
        if ft == "boolean":
            return _to_bool(text)
        if ft == "integer":
            return int(text)
        if ft == "float":
            return float(text)
        if ft == "enum":
            return text
        if ft == "list":
            return [p for p in text.split(",") if p != ""] if text else []
    except ValueError as e:
        raise FlagError("flag --%s=%s: %s" % (name, text, e))
    raise FlagError("unknown flag type %s" % ft)


def validate_value(name: str, value: Any) -> None:
    """Bounds / enum checks (T/benchmark_cnn.py:962-990)."""
    spec = param_specs[name]
    if value is None:
        return
    kw = spec.kwargs
    if "lower_bound" in kw and value < kw["lower_bound"]:
        raise ValueError("Param %s value of %s is lower than the lower bound of %s" %
                         (name, value, kw["lower_bound"]))
    if "upper_bound" in kw and value > kw["upper_bound"]:
        raise ValueError("Param %s value of %s is higher than the upper bound of %s" %
                         (name, value, kw["upper_bound"]))
    if "enum_values" in kw and value not in kw["enum_values"]:
        raise ValueError("Param %s of value %s is not in %s" % (name, value, kw["enum_values"]))


def parse_argv(argv: Sequence[str]) -> Dict[str, Any]:
    """Parses absl-style arguments into {name: value} for the flags that were given."""
    out: Dict[str, Any] = {}
    args = list(argv)
    i = 0
    while i < len(args):
        a = args[i]
        i += 1
        if a == "--":
            if i < len(args):
                raise FlagError("Received unknown positional arguments: %s" % args[i:])
            break
        if not a.startswith("-"):
            raise FlagError("Received unknown positional arguments: %s" % [a])
        body = a.lstrip("-")
        if "=" in body:
            name, value = body.split("=", 1)
            has_value = True
        else:
            name, value, has_value = body, None, False
        name = aliases.get(name, name)
        if name not in param_specs:
            if name.startswith("no") and aliases.get(name[2:], name[2:]) in param_specs and \
                    param_specs[aliases.get(name[2:], name[2:])].flag_type == "boolean" and \
                    not has_value:
                out[aliases.get(name[2:], name[2:])] = False
                continue
            raise FlagError("Unknown command line flag '%s'" % body.split("=")[0])
        spec = param_specs[name]
        if spec.flag_type == "boolean":
            out[name] = _to_bool(value) if has_value else True
            continue
        if not has_value:
            if i >= len(args):
                raise FlagError("Missing value for flag --%s" % name)
            value = args[i]
            i += 1
        out[name] = convert(name, value)
    return out


def help_text() -> str:
    lines = []
    for name, spec in param_specs.items():
        extra = ""
        if "enum_values" in spec.kwargs:
            extra = " <%s>" % "|".join(str(v) for v in spec.kwargs["enum_values"])
        lines.append("  --%s%s: %s (default: %r)" % (name, extra, spec.description,
                                                    spec.default_value))
    return "\n".join(lines)
