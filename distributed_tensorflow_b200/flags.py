"""Command-line flags with the semantics the reference relies on from ``tf.app.flags``.

The reference declares flags with ``flags.DEFINE_integer/string/float`` and reads ``FLAGS.x``
without ever calling ``tf.app.run()`` (/root/reference/trainer/task.py:17-31, 310-311), so parsing
is lazy (first attribute access), runs against ``sys.argv`` and silently ignores flags it does not
know.  Both ``--name=value`` and ``--name value`` are accepted; booleans also accept ``--name`` /
``--noname``.
"""
from __future__ import annotations

import sys
from typing import Any, Callable


class _Flag:
    __slots__ = ("name", "default", "help", "parser", "is_bool")

    def __init__(self, name: str, default: Any, help: str, parser: Callable[[str], Any], is_bool: bool = False):
        self.name, self.default, self.help, self.parser, self.is_bool = name, default, help, parser, is_bool


def _parse_bool(s: str) -> bool:
    v = s.strip().lower()
    if v in ("1", "true", "t", "yes", "y"):
        return True
    if v in ("0", "false", "f", "no", "n"):
        return False
    raise ValueError("not a boolean: %r" % s)


class FlagValues:
    def __init__(self):
        object.__setattr__(self, "_defs", {})
        object.__setattr__(self, "_values", {})
        object.__setattr__(self, "_parsed", False)

    # -- definition ------------------------------------------------------------------------
    def _define(self, flag: _Flag) -> None:
        self._defs[flag.name] = flag
        self._values[flag.name] = flag.default
        object.__setattr__(self, "_parsed", False)

    # -- parsing ---------------------------------------------------------------------------
    def _parse(self, argv=None) -> list[str]:
        """Parse ``argv`` (default ``sys.argv[1:]``); returns the arguments that were not flags."""
        args = list(sys.argv[1:] if argv is None else argv)
        rest: list[str] = []
        i = 0
        while i < len(args):
            a = args[i]
            i += 1
            if a == "--":
                rest.extend(args[i:])
                break
            if not a.startswith("-") or a in ("-", "--"):
                rest.append(a)
                continue
            body = a.lstrip("-")
            name, eq, val = body.partition("=")
            name = name.replace("-", "_") if name not in self._defs else name
            flag = self._defs.get(name)
            if flag is None and name.startswith("no") and name[2:] in self._defs and self._defs[name[2:]].is_bool:
                self._values[name[2:]] = False
                continue
            if flag is None:
                # unknown flag: tolerated, like tf.app.flags when main() is called directly.
                # A following bare token is left in `rest` rather than swallowed.
                continue
            if flag.is_bool and not eq:
                self._values[name] = True
                continue
            if not eq:
                if i >= len(args):
                    raise ValueError("flag --%s needs a value" % name)
                val = args[i]
                i += 1
            try:
                self._values[name] = flag.parser(val)
            except ValueError as e:
                raise ValueError("flag --%s=%s: %s" % (name, val, e)) from None
        object.__setattr__(self, "_parsed", True)
        return rest

    def __call__(self, argv=None):
        return self._parse(argv)

    def __getattr__(self, name: str) -> Any:
        defs = object.__getattribute__(self, "_defs")
        if name not in defs:
            raise AttributeError(name)
        if not object.__getattribute__(self, "_parsed"):
            self._parse()
        return object.__getattribute__(self, "_values")[name]

    def __setattr__(self, name: str, value: Any) -> None:
        if name not in self._defs:
            raise AttributeError(name)
        self._values[name] = value

    def __contains__(self, name: str) -> bool:
        return name in self._defs

    def flag_values_dict(self) -> dict:
        if not self._parsed:
            self._parse()
        return dict(self._values)

    def reset(self) -> None:
        """Forget parsed values (keeps definitions); the next access re-parses ``sys.argv``."""
        for k, f in self._defs.items():
            self._values[k] = f.default
        object.__setattr__(self, "_parsed", False)


FLAGS = FlagValues()


def DEFINE_integer(name: str, default, help: str = "", flag_values: FlagValues = None) -> None:
    (flag_values or FLAGS)._define(_Flag(name, default, help, int))


def DEFINE_float(name: str, default, help: str = "", flag_values: FlagValues = None) -> None:
    (flag_values or FLAGS)._define(_Flag(name, default, help, float))


def DEFINE_string(name: str, default, help: str = "", flag_values: FlagValues = None) -> None:
    (flag_values or FLAGS)._define(_Flag(name, default, help, str))


def DEFINE_boolean(name: str, default, help: str = "", flag_values: FlagValues = None) -> None:
    (flag_values or FLAGS)._define(_Flag(name, default, help, _parse_bool, is_bool=True))


DEFINE_bool = DEFINE_boolean
