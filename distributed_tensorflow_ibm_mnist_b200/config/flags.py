"""TensorFlow-style command line flags (``tf.app.flags`` stand-in).

The reference defines its flags at import time and reads them from a global ``FLAGS`` object
(main.py:10-32, inference.py:14-30; mnist_input.py:9 shares the same object).  This module keeps that contract —
``DEFINE_string/integer/float/boolean``, a process-wide ``FLAGS``, ``--name=value`` / ``--name value`` /
``--flag`` / ``--noflag`` syntax — and adds aliases (``--task_index`` for ``--task_id``).
"""
from __future__ import annotations

import sys
from typing import Any, Callable


class FlagError(ValueError):
    pass


def _to_bool(s: str) -> bool:
    v = s.strip().lower()
    if v in ("1", "true", "t", "yes", "y"):
        return True
    if v in ("0", "false", "f", "no", "n"):
        return False
    raise FlagError(f"not a boolean: {s!r}")


class _FlagValues:
    def __init__(self):
        object.__setattr__(self, "_defs", {})      # name -> (parser, default, help)
        object.__setattr__(self, "_values", {})
        object.__setattr__(self, "_aliases", {})
        object.__setattr__(self, "_parsed", False)

    # definition --------------------------------------------------------------------------------------
    def _define(self, name: str, default: Any, help_: str, parser: Callable[[str], Any]) -> None:
        self._defs[name] = (parser, default, help_)
        self._values.setdefault(name, default)

    def alias(self, alias: str, target: str) -> None:
        self._aliases[alias] = target

    # access ------------------------------------------------------------------------------------------
    def __getattr__(self, name: str) -> Any:
        name = self._aliases.get(name, name)
        if name in self._values:
            return self._values[name]
        raise AttributeError(f"unknown flag {name!r}")

    def __setattr__(self, name: str, value: Any) -> None:
        name = self._aliases.get(name, name)
        if name not in self._defs:
            raise AttributeError(f"unknown flag {name!r}")
        self._values[name] = value

    def __contains__(self, name: str) -> bool:
        return self._aliases.get(name, name) in self._defs

    def flag_values_dict(self) -> dict:
        return dict(self._values)

    # parsing -----------------------------------------------------------------------------------------
    def parse(self, argv: list[str] | None = None, *, known_only: bool = False) -> list[str]:
        """Parse ``argv`` (without the program name); returns the positional / unknown remainder."""
        argv = list(sys.argv[1:] if argv is None else argv)
        rest: list[str] = []
        i = 0
        while i < len(argv):
            tok = argv[i]
            i += 1
            if tok == "--":
                rest += argv[i:]
                break
            if not tok.startswith("-") or tok == "-":
                rest.append(tok)
                continue
            body = tok.lstrip("-")
            name, eq, val = body.partition("=")
            name = self._aliases.get(name, name)
            if name not in self._defs:
                if name.startswith("no") and self._aliases.get(name[2:], name[2:]) in self._defs and not eq:
                    target = self._aliases.get(name[2:], name[2:])
                    if self._defs[target][0] is _to_bool:
                        self._values[target] = False
                        continue
                if known_only:
                    rest.append(tok)
                    continue
                raise FlagError(f"unknown flag --{name}")
            parser = self._defs[name][0]
            if not eq:
                if parser is _to_bool:
                    # `--flag` alone means true; `--flag false` is also accepted
                    if i < len(argv) and argv[i].lower() in ("true", "false", "0", "1"):
                        val = argv[i]
                        i += 1
                    else:
                        val = "true"
                else:
                    if i >= len(argv):
                        raise FlagError(f"flag --{name} needs a value")
                    val = argv[i]
                    i += 1
            try:
                self._values[name] = parser(val)
            except (ValueError, FlagError) as e:
                raise FlagError(f"bad value for --{name}: {val!r} ({e})") from None
        object.__setattr__(self, "_parsed", True)
        return rest

    def reset(self) -> None:
        for name, (_, default, _) in self._defs.items():
            self._values[name] = default

    def usage(self) -> str:
        lines = []
        for name, (parser, default, help_) in sorted(self._defs.items()):
            lines.append(f"  --{name}={default!r}\n      {help_}")
        return "\n".join(lines)


FLAGS = _FlagValues()


def DEFINE_string(name: str, default: str | None, help: str = "") -> None:  # noqa: N802 (TF naming)
    FLAGS._define(name, default, help, str)


def DEFINE_integer(name: str, default: int | None, help: str = "") -> None:  # noqa: N802
    FLAGS._define(name, default, help, int)


def DEFINE_float(name: str, default: float | None, help: str = "") -> None:  # noqa: N802
    FLAGS._define(name, default, help, float)


def DEFINE_boolean(name: str, default: bool, help: str = "") -> None:  # noqa: N802
    FLAGS._define(name, default, help, _to_bool)


DEFINE_bool = DEFINE_boolean


def DEFINE_alias(alias: str, target: str) -> None:  # noqa: N802
    FLAGS.alias(alias, target)


def run(main: Callable, argv: list[str] | None = None) -> None:
    """``tf.app.run`` stand-in: parse flags, call ``main(remaining_argv)``, exit with its return value."""
    args = list(sys.argv[1:] if argv is None else argv)
    flag_part = args[:args.index("--")] if "--" in args else args
    if any(a in ("--help", "-h", "-help", "--helpshort", "--helpfull") for a in flag_part):
        # tf.app.run prints the program's flags and exits (status 1 in TF-1.x; 0 here, like every other tool)
        doc = (getattr(sys.modules.get("__main__"), "__doc__", None) or "").strip()
        print((doc + "\n\n" if doc else "") + f"flags:\n{FLAGS.usage()}")
        sys.exit(0)
    try:
        rest = FLAGS.parse(argv)
    except FlagError as e:
        print(f"flag error: {e}\n\nflags:\n{FLAGS.usage()}", file=sys.stderr)
        sys.exit(2)
    sys.exit(main([sys.argv[0]] + rest))
