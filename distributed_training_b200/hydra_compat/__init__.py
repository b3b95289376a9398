"""Minimal Hydra-compatible configuration layer.

The reference is configured through Hydra 1.3 (``@hydra.main`` at
/root/reference/src/distributed_trainer.py:243, config tree at
/root/reference/conf/config.yaml:1-14).  hydra-core / omegaconf are not part of
the B200 image and there is no network, so this module implements exactly the
subset the trainer relies on and transparently defers to real Hydra when it is
importable:

* ``main(version_base, config_path, config_name)`` decorator
* defaults-list composition with config groups and ``_self_``
* command line overrides: ``a.b=value``, ``+a.b=value``, ``~a.b``, ``group=option``
* interpolations ``${a.b}``, ``${now:%fmt}``, ``${hydra:runtime.output_dir}``,
  ``${oc.env:VAR,default}``
* ``hydra.run.dir`` creation, ``hydra.job.chdir`` and the ``.hydra/`` dump
  (``config.yaml``, ``overrides.yaml``, ``hydra.yaml``)
* attribute / item / ``.get`` access on the resulting config object

``install_stub_modules()`` registers ``hydra`` and ``omegaconf`` stand-ins in
``sys.modules`` so that code written against the real packages (the unmodified
reference used by the baseline arm of ``bench.py``) imports and runs.
"""
from __future__ import annotations

import copy
import datetime as _dt
import functools
import os
import re
import sys
import types
from pathlib import Path
from typing import Any, Callable, Dict, Iterator, List, Optional, Sequence, Tuple

import yaml

__all__ = [
    "ConfigNode",
    "ConfigError",
    "compose",
    "main",
    "to_container",
    "to_yaml",
    "install_stub_modules",
    "using_real_hydra",
]


class ConfigError(Exception):
    """Raised for malformed configs, overrides or interpolations."""


_MISSING = object()


class ConfigNode:
    """Dict-like config node with attribute access and lazy interpolation.

    Mirrors the parts of ``omegaconf.DictConfig`` the trainer uses:
    ``cfg.train.batch_size``, ``cfg.train.get("device", "auto")``,
    ``cfg["train"]``, ``"x" in cfg``, iteration over keys.
    """

    __slots__ = ("_data", "_root", "_ctx")

    def __init__(self, data: Dict[str, Any], root: "Optional[ConfigNode]" = None,
                 ctx: Optional[Dict[str, Any]] = None) -> None:
        object.__setattr__(self, "_data", data)
        object.__setattr__(self, "_root", root if root is not None else self)
        object.__setattr__(self, "_ctx", ctx if ctx is not None else {})

    # -- access -----------------------------------------------------------
    def _wrap(self, value: Any) -> Any:
        if isinstance(value, dict):
            return ConfigNode(value, self._root, self._ctx)
        if isinstance(value, list):
            return [self._wrap(v) for v in value]
        if isinstance(value, str) and "${" in value:
            return _resolve_string(value, self._root, self._ctx)
        return value

    def __getattr__(self, name: str) -> Any:
        if name.startswith("__"):
            raise AttributeError(name)
        try:
            return self._wrap(self._data[name])
        except KeyError:
            raise AttributeError(
                f"Missing key {name!r}; available: {sorted(self._data)}"
            ) from None

    def __getitem__(self, name: str) -> Any:
        return self._wrap(self._data[name])

    def __setattr__(self, name: str, value: Any) -> None:
        self._data[name] = _unwrap(value)

    def __setitem__(self, name: str, value: Any) -> None:
        self._data[name] = _unwrap(value)

    def __contains__(self, name: object) -> bool:
        return name in self._data

    def __iter__(self) -> Iterator[str]:
        return iter(self._data)

    def __len__(self) -> int:
        return len(self._data)

    def get(self, name: str, default: Any = None) -> Any:
        if name in self._data and self._data[name] is not None:
            return self._wrap(self._data[name])
        return default

    def keys(self):
        return self._data.keys()

    def items(self):
        return [(k, self._wrap(v)) for k, v in self._data.items()]

    def values(self):
        return [self._wrap(v) for v in self._data.values()]

    def __repr__(self) -> str:
        return f"ConfigNode({to_container(self, resolve=False)!r})"

    def __eq__(self, other: object) -> bool:
        if isinstance(other, ConfigNode):
            return to_container(self) == to_container(other)
        if isinstance(other, dict):
            return to_container(self) == other
        return NotImplemented


def _unwrap(value: Any) -> Any:
    if isinstance(value, ConfigNode):
        return value._data
    return value


# ---------------------------------------------------------------------------
# interpolation
# ---------------------------------------------------------------------------
_INTERP = re.compile(r"\$\{([^${}]+)\}")


def _select(root: ConfigNode, dotted: str) -> Any:
    node: Any = root._data
    for part in dotted.split("."):
        if isinstance(node, dict) and part in node:
            node = node[part]
        else:
            raise ConfigError(f"Interpolation key not found: {dotted!r}")
    return node


def _resolve_token(token: str, root: ConfigNode, ctx: Dict[str, Any]) -> Any:
    token = token.strip()
    if ":" in token:
        resolver, _, arg = token.partition(":")
        resolver = resolver.strip()
        if resolver == "now":
            now = ctx.setdefault("now", _dt.datetime.now())
            return now.strftime(arg)
        if resolver == "hydra":
            value = _select(root, "hydra." + arg.strip())
            if isinstance(value, str) and "${" in value:
                value = _resolve_string(value, root, ctx)
            return value
        if resolver == "oc.env":
            var, _, default = arg.partition(",")
            if var.strip() in os.environ:
                return os.environ[var.strip()]
            if default == "":
                raise ConfigError(f"Environment variable {var!r} not set")
            return default.strip()
        raise ConfigError(f"Unsupported resolver {resolver!r} in ${{{token}}}")
    value = _select(root, token)
    if isinstance(value, str) and "${" in value:
        value = _resolve_string(value, root, ctx)
    return value


def _resolve_string(text: str, root: ConfigNode, ctx: Dict[str, Any], depth: int = 0) -> Any:
    if depth > 16:
        raise ConfigError(f"Interpolation recursion too deep in {text!r}")
    whole = _INTERP.fullmatch(text)
    if whole:  # keep the referenced value's type
        value = _resolve_token(whole.group(1), root, ctx)
        if isinstance(value, dict):
            return ConfigNode(value, root, ctx)
        return value
    out = _INTERP.sub(lambda m: str(_resolve_token(m.group(1), root, ctx)), text)
    if "${" in out and out != text:
        return _resolve_string(out, root, ctx, depth + 1)
    return out


def to_container(cfg: Any, resolve: bool = True) -> Any:
    """Plain ``dict``/``list`` copy of a config (``OmegaConf.to_container``)."""
    if isinstance(cfg, ConfigNode):
        if not resolve:
            return copy.deepcopy(cfg._data)
        return {k: to_container(cfg[k], True) for k in cfg._data}
    if isinstance(cfg, list):
        return [to_container(v, resolve) for v in cfg]
    return cfg


def to_yaml(cfg: Any, resolve: bool = False) -> str:
    return yaml.safe_dump(to_container(cfg, resolve), sort_keys=False, default_flow_style=False)


# ---------------------------------------------------------------------------
# composition
# ---------------------------------------------------------------------------
def _load_yaml(path: Path) -> Dict[str, Any]:
    if not path.exists():
        raise ConfigError(f"Config file not found: {path}")
    with open(path, "r", encoding="utf-8") as fh:
        data = yaml.safe_load(fh)
    if data is None:
        return {}
    if not isinstance(data, dict):
        raise ConfigError(f"Top level of {path} must be a mapping")
    return data


def _deep_merge(dst: Dict[str, Any], src: Dict[str, Any]) -> Dict[str, Any]:
    for key, value in src.items():
        if isinstance(value, dict) and isinstance(dst.get(key), dict):
            _deep_merge(dst[key], value)
        else:
            dst[key] = copy.deepcopy(value)
    return dst


def _parse_value(text: str) -> Any:
    if text == "":
        return ""
    try:
        value = yaml.safe_load(text)
    except yaml.YAMLError:
        return text
    if isinstance(value, str):  # YAML 1.1 misses floats such as "1e-2"; Hydra's grammar accepts them
        try:
            return int(value)
        except ValueError:
            try:
                return float(value)
            except ValueError:
                return value
    return value


def _set_dotted(data: Dict[str, Any], dotted: str, value: Any, *, must_exist: bool) -> None:
    parts = dotted.split(".")
    node = data
    for part in parts[:-1]:
        if part not in node or not isinstance(node[part], dict):
            if must_exist and not dotted.startswith("hydra."):
                raise ConfigError(
                    f"Could not override {dotted!r}: key {part!r} is not in the config "
                    f"(use +{dotted}=... to add it)"
                )
            node[part] = {}
        node = node[part]
    if must_exist and parts[-1] not in node and not dotted.startswith("hydra."):
        raise ConfigError(
            f"Could not override {dotted!r}: no such key (use +{dotted}=... to add it)"
        )
    node[parts[-1]] = value


def _del_dotted(data: Dict[str, Any], dotted: str) -> None:
    parts = dotted.split(".")
    node = data
    for part in parts[:-1]:
        node = node.get(part, {})
    node.pop(parts[-1], None)


_HYDRA_DEFAULTS: Dict[str, Any] = {
    "run": {"dir": "outputs/${now:%Y-%m-%d}/${now:%H-%M-%S}"},
    "job": {"chdir": None, "name": "app"},
    "output_subdir": ".hydra",
    "runtime": {"output_dir": "${hydra.run.dir}", "cwd": None},
}


def _split_overrides(config_dir: Path, overrides: Sequence[str]) -> Tuple[Dict[str, str], List[Tuple[str, str, Any]]]:
    """Separate ``group=option`` choices from value overrides."""
    groups: Dict[str, str] = {}
    values: List[Tuple[str, str, Any]] = []
    for raw in overrides:
        if raw.startswith("~"):
            values.append(("del", raw[1:].split("=", 1)[0], None))
            continue
        if "=" not in raw:
            raise ConfigError(f"Override {raw!r} is not of the form key=value")
        key, _, text = raw.partition("=")
        mode = "set"
        if key.startswith("++"):
            key, mode = key[2:], "force"
        elif key.startswith("+"):
            key, mode = key[1:], "add"
        group_dir = config_dir / key.replace(".", "/")
        if mode == "set" and group_dir.is_dir() and "." not in key:
            groups[key] = text
        else:
            values.append((mode, key, _parse_value(text)))
    return groups, values


def compose(config_dir: os.PathLike, config_name: str = "config",
            overrides: Sequence[str] = ()) -> ConfigNode:
    """Compose ``<config_dir>/<config_name>.yaml`` with its defaults list.

    Equivalent of ``hydra.compose`` for the feature subset in the module
    docstring; group files are merged under their group key, ``_self_`` marks
    where the primary file's own content lands.
    """
    config_dir = Path(config_dir).resolve()
    primary = _load_yaml(config_dir / f"{config_name}.yaml")
    defaults = primary.pop("defaults", [])
    group_choice, value_overrides = _split_overrides(config_dir, overrides)

    merged: Dict[str, Any] = {}
    choices: Dict[str, str] = {}
    saw_self = False
    for entry in list(defaults) + ([] if "_self_" in defaults else ["_self_"]):
        if entry == "_self_":
            _deep_merge(merged, primary)
            saw_self = True
            continue
        if isinstance(entry, str):
            _deep_merge(merged, _load_yaml(config_dir / f"{entry}.yaml"))
            continue
        if not isinstance(entry, dict) or len(entry) != 1:
            raise ConfigError(f"Unsupported defaults entry: {entry!r}")
        (group, option), = entry.items()
        group = str(group)
        optional = group.startswith("optional ")
        if optional:
            group = group[len("optional "):]
        option = group_choice.pop(group, option)
        if option is None or option == "null":
            continue
        path = config_dir / group / f"{option}.yaml"
        if not path.exists():
            if optional:
                continue
            avail = sorted(p.stem for p in (config_dir / group).glob("*.yaml"))
            raise ConfigError(
                f"Could not find '{group}/{option}'. Available options in '{group}': {avail}"
            )
        choices[group] = str(option)
        node = merged
        for part in group.split("/")[:-1]:
            node = node.setdefault(part, {})
        leaf = group.split("/")[-1]
        node[leaf] = _deep_merge(node.get(leaf, {}) if isinstance(node.get(leaf), dict) else {},
                                 _load_yaml(path))
    assert saw_self
    # group overrides for groups not in the defaults list (e.g. `+model=x` style)
    for group, option in group_choice.items():
        path = config_dir / group / f"{option}.yaml"
        if not path.exists():
            avail = sorted(p.stem for p in (config_dir / group).glob("*.yaml"))
            raise ConfigError(
                f"Could not find '{group}/{option}'. Available options in '{group}': {avail}"
            )
        choices[group] = str(option)
        merged[group] = _deep_merge(merged.get(group, {}) if isinstance(merged.get(group), dict) else {},
                                    _load_yaml(path))

    hydra_node = _deep_merge(copy.deepcopy(_HYDRA_DEFAULTS), merged.get("hydra", {}) or {})
    merged["hydra"] = hydra_node
    hydra_node.setdefault("runtime", {})["choices"] = choices
    hydra_node["overrides"] = {"task": list(overrides)}

    for mode, key, value in value_overrides:
        if mode == "del":
            _del_dotted(merged, key)
        else:
            _set_dotted(merged, key, value, must_exist=(mode == "set"))
    return ConfigNode(merged)


# ---------------------------------------------------------------------------
# @main
# ---------------------------------------------------------------------------
def using_real_hydra() -> bool:
    """True when the genuine hydra-core package is importable."""
    mod = sys.modules.get("hydra")
    if mod is not None and getattr(mod, "__dtb200_stub__", False):
        return False
    try:
        import importlib.util

        return importlib.util.find_spec("hydra") is not None
    except (ImportError, ValueError):
        return False


def _strip_hydra(cfg: ConfigNode) -> ConfigNode:
    """User-facing config: everything but the ``hydra`` node, resolved against the full tree."""
    data = {k: v for k, v in cfg._data.items() if k != "hydra"}
    view = ConfigNode(data, root=cfg._root, ctx=cfg._ctx)
    return view


def run_job(task: Callable[[Any], Any], config_dir: os.PathLike, config_name: str,
            overrides: Sequence[str]) -> Any:
    """Compose, create the run dir, dump ``.hydra/`` and call ``task(cfg)``."""
    cfg = compose(config_dir, config_name, overrides)
    original_cwd = os.getcwd()
    cfg._data["hydra"]["runtime"]["cwd"] = original_cwd
    run_dir = Path(str(cfg.hydra.run.dir))
    if not run_dir.is_absolute():
        run_dir = Path(original_cwd) / run_dir
    cfg._data["hydra"]["run"]["dir"] = str(run_dir)
    cfg._data["hydra"]["runtime"]["output_dir"] = str(run_dir)
    run_dir.mkdir(parents=True, exist_ok=True)

    user_cfg = _strip_hydra(cfg)
    subdir = cfg.hydra.get("output_subdir", ".hydra")
    # Every rank of a torchrun job composes the same config; only one needs to dump it.
    if subdir and int(os.environ.get("RANK", "0")) == 0:
        meta = run_dir / str(subdir)
        meta.mkdir(parents=True, exist_ok=True)
        (meta / "config.yaml").write_text(to_yaml(user_cfg, resolve=False))
        (meta / "overrides.yaml").write_text(yaml.safe_dump(list(overrides)))
        (meta / "hydra.yaml").write_text(
            yaml.safe_dump({"hydra": to_container(cfg.hydra, resolve=True)}, sort_keys=False)
        )
    chdir = cfg.hydra.job.get("chdir", None)
    try:
        if chdir:
            os.chdir(run_dir)
        return task(user_cfg)
    finally:
        if chdir:
            os.chdir(original_cwd)


def main(version_base: Optional[str] = None, config_path: Optional[str] = None,
         config_name: Optional[str] = "config") -> Callable[[Callable[[Any], Any]], Callable[..., Any]]:
    """Drop-in for ``hydra.main``.

    ``config_path`` is relative to the file that defines the decorated
    function, as in Hydra.  Calling the decorated function with an explicit
    config (``fn(cfg)``) bypasses composition, which is what Hydra does too.
    """

    def decorator(task: Callable[[Any], Any]) -> Callable[..., Any]:
        @functools.wraps(task)
        def wrapper(cfg: Any = _MISSING, *, argv: Optional[Sequence[str]] = None) -> Any:
            if cfg is not _MISSING:
                return task(cfg)
            src_file = sys.modules[task.__module__].__file__ if task.__module__ in sys.modules else None
            base = Path(src_file).resolve().parent if src_file else Path.cwd()
            cfg_dir = (base / (config_path or ".")).resolve()
            args = list(sys.argv[1:] if argv is None else argv)
            overrides = [a for a in args if not a.startswith("--")]
            for flag in (a for a in args if a.startswith("--")):
                if flag in ("--cfg", "--help", "-h"):
                    print(to_yaml(_strip_hydra(compose(cfg_dir, config_name or "config", overrides))))
                    return None
            return run_job(task, cfg_dir, config_name or "config", overrides)

        return wrapper

    return decorator


# ---------------------------------------------------------------------------
# stub modules for third-party code that imports hydra / omegaconf directly
# ---------------------------------------------------------------------------
def install_stub_modules(force: bool = False) -> bool:
    """Expose this shim as ``hydra`` and ``omegaconf`` in ``sys.modules``.

    No-op (returns False) when the real packages are importable, unless
    ``force``.  Used by ``baseline/`` to run the unmodified reference.
    """
    if not force and using_real_hydra():
        return False
    hydra_mod = types.ModuleType("hydra")
    hydra_mod.__dtb200_stub__ = True  # type: ignore[attr-defined]
    hydra_mod.main = main  # type: ignore[attr-defined]
    hydra_mod.compose = compose  # type: ignore[attr-defined]

    oc_mod = types.ModuleType("omegaconf")
    oc_mod.__dtb200_stub__ = True  # type: ignore[attr-defined]
    oc_mod.DictConfig = ConfigNode  # type: ignore[attr-defined]

    class _OmegaConf:
        @staticmethod
        def to_container(cfg: Any, resolve: bool = False) -> Any:
            return to_container(cfg, resolve)

        @staticmethod
        def to_yaml(cfg: Any, resolve: bool = False) -> str:
            return to_yaml(cfg, resolve)

        @staticmethod
        def create(obj: Optional[Dict[str, Any]] = None) -> ConfigNode:
            return ConfigNode(copy.deepcopy(obj or {}))

    oc_mod.OmegaConf = _OmegaConf  # type: ignore[attr-defined]
    sys.modules["hydra"] = hydra_mod
    sys.modules["omegaconf"] = oc_mod
    return True
