"""Mini-Hydra: just enough of Hydra 1.3 / OmegaConf to compose the reference's ``config/*.yaml`` (kept verbatim under
``learningsparsepreconditioner4gpu_b200/config/``) the way ``@hydra.main(config_path="config", config_name="basic")``
does for ``train.py:28`` / ``infer.py:166`` -- hydra-core and omegaconf are not installed here.

Supported (what the reference's YAML files and launch scripts use, training/*.sh, misc/*.sh):

* a ``defaults:`` list of sibling file names and ``_self_`` (config/basic.yaml:1-6), merged in list order;
* ``${a.b}`` interpolation, resolved against the ROOT after all overrides (config/gnn.yaml, basic.yaml:11);
* ``???`` mandatory values: reading one that was never overridden raises ``MissingMandatoryValue``;
* command-line overrides ``a.b=v`` (key must exist), ``+a.b=v`` (key must be new), ``++a.b=v`` (either), ``~a.b``
  (delete), values parsed as YAML; ``--config-name=NAME`` / ``-cn NAME``, ``--config-path=DIR`` / ``-cp DIR``;
* attribute and item access, ``cfg.get(key, default)``, ``in``, ``**cfg`` splatting, ``to_container``.
"""
from __future__ import annotations

import os
import re
from typing import Any, Dict, Iterator, List, Optional, Sequence, Tuple

import yaml

CONFIG_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "config")
_INTERP = re.compile(r"\$\{([^${}]+)\}")
MISSING = "???"


class ConfigError(Exception):
    pass


class MissingMandatoryValue(ConfigError):
    pass


class ConfigKeyError(ConfigError, KeyError):
    pass


def _deep_merge(dst: dict, src: dict) -> dict:
    for k, v in src.items():
        if isinstance(v, dict) and isinstance(dst.get(k), dict):
            _deep_merge(dst[k], v)
        else:
            dst[k] = _copy(v)
    return dst


def _copy(v):
    if isinstance(v, dict):
        return {k: _copy(x) for k, x in v.items()}
    if isinstance(v, list):
        return [_copy(x) for x in v]
    return v


class _Loader(yaml.SafeLoader):
    """YAML 1.1 reads ``1e-3`` (no dot) as a string; OmegaConf reads it as a float (config/trainer.yaml:9)."""


_Loader.add_implicit_resolver(
    "tag:yaml.org,2002:float",
    re.compile(r"^[-+]?(?:[0-9][0-9_]*\.[0-9_]*(?:[eE][-+]?[0-9]+)?|\.[0-9_]+(?:[eE][-+]?[0-9]+)?|[0-9][0-9_]*[eE][-+]?[0-9]+"
               r"|\.(?:inf|Inf|INF)|\.(?:nan|NaN|NAN))$"),
    list("-+0123456789."))


def _parse(text: str):
    return yaml.load(text, Loader=_Loader)


def _load_file(config_dir: str, name: str) -> dict:
    path = os.path.join(config_dir, name if name.endswith((".yaml", ".yml")) else name + ".yaml")
    if not os.path.exists(path):
        raise ConfigError(f"config file not found: {path}")
    with open(path) as f:
        return _parse(f.read()) or {}


def _compose_file(config_dir: str, name: str) -> dict:
    raw = _load_file(config_dir, name)
    defaults = raw.pop("defaults", None)
    if defaults is None:
        return raw
    out: dict = {}
    seen_self = False
    for entry in defaults:
        if entry == "_self_":
            _deep_merge(out, raw)
            seen_self = True
        elif isinstance(entry, str):
            _deep_merge(out, _compose_file(config_dir, entry))
        elif isinstance(entry, dict):  # "group: option" -> config_dir/group/option.yaml placed under key `group`
            for group, option in entry.items():
                if option is None:
                    continue
                _deep_merge(out.setdefault(group, {}), _compose_file(os.path.join(config_dir, group), option))
        else:
            raise ConfigError(f"unsupported defaults entry {entry!r} in {name}")
    if not seen_self:
        _deep_merge(out, raw)  # Hydra 1.1+: _self_ is last when it is not listed
    return out


def _walk(root: dict, dotted: str, create: bool = False) -> Tuple[dict, str]:
    parts = dotted.split(".")
    node = root
    for p in parts[:-1]:
        if p not in node or not isinstance(node[p], dict):
            if not create:
                raise ConfigKeyError(f"key {dotted!r}: {p!r} is not a config group")
            node[p] = {}
        node = node[p]
    return node, parts[-1]


def apply_overrides(root: dict, overrides: Sequence[str]) -> dict:
    for ov in overrides:
        if ov.startswith("~"):
            node, leaf = _walk(root, ov[1:].split("=")[0])
            if leaf not in node:
                raise ConfigKeyError(f"cannot delete {ov[1:]!r}: no such key")
            del node[leaf]
            continue
        if "=" not in ov:
            raise ConfigError(f"override {ov!r} is not of the form key=value")
        key, text = ov.split("=", 1)
        force = key.startswith("++")
        add = key.startswith("+") and not force
        key = key.lstrip("+")
        value = _parse(text) if text != "" else ""
        node, leaf = _walk(root, key, create=add or force)
        if add and leaf in node:
            raise ConfigError(f"could not append {key!r}: the key is already in the config (use {key}=... or ++{key}=...)")
        if not add and not force and leaf not in node:
            raise ConfigKeyError(f"could not override {key!r}: no such key in the config (use +{key}=... to add it)")
        node[leaf] = value
    return root


class Cfg:
    """Attribute-access view of a composed config (the part of ``DictConfig`` the reference touches)."""

    __slots__ = ("_data", "_root", "_path")

    def __init__(self, data: dict, root: Optional[dict] = None, path: str = ""):
        object.__setattr__(self, "_data", data)
        object.__setattr__(self, "_root", data if root is None else root)
        object.__setattr__(self, "_path", path)

    # ---- resolution ------------------------------------------------------------------------------------
    def _lookup(self, dotted: str, depth: int):
        node: Any = self._root
        for p in dotted.strip().split("."):
            if not isinstance(node, dict) or p not in node:
                raise ConfigKeyError(f"interpolation ${{{dotted}}} not found")
            node = node[p]
        return self._resolve(node, dotted, depth + 1)

    def _resolve(self, v, where: str, depth: int = 0):
        if depth > 32:
            raise ConfigError(f"interpolation cycle at {where!r}")
        if isinstance(v, str):
            if v == MISSING:
                raise MissingMandatoryValue(f"missing mandatory value: {where}")
            m = _INTERP.fullmatch(v)
            if m:  # whole-value interpolation keeps the type of its target
                return self._lookup(m.group(1), depth)
            return _INTERP.sub(lambda mm: str(self._lookup(mm.group(1), depth)), v)
        return v

    def _wrap(self, key: str, v):
        where = f"{self._path}.{key}" if self._path else key
        v = self._resolve(v, where)
        if isinstance(v, dict):
            return Cfg(v, self._root, where)
        if isinstance(v, list):
            return [Cfg(x, self._root, where) if isinstance(x, dict) else self._resolve(x, where) for x in v]
        return v

    # ---- mapping protocol ------------------------------------------------------------------------------
    def __getattr__(self, key: str):
        if key.startswith("__"):
            raise AttributeError(key)
        try:
            return self._wrap(key, self._data[key])
        except KeyError:
            raise AttributeError(f"config has no key {(self._path + '.' if self._path else '') + key!r}") from None

    def __getitem__(self, key: str):
        if key not in self._data:
            raise ConfigKeyError(key)
        return self._wrap(key, self._data[key])

    def __setattr__(self, key: str, value):
        self._data[key] = value

    __setitem__ = __setattr__

    def get(self, key: str, default=None):
        if key not in self._data or self._data[key] is None:
            return default
        return self._wrap(key, self._data[key])

    def __contains__(self, key) -> bool:
        return key in self._data

    def keys(self):
        return self._data.keys()

    def __iter__(self) -> Iterator[str]:
        return iter(self._data)

    def __len__(self) -> int:
        return len(self._data)

    def items(self):
        return [(k, self[k]) for k in self._data]

    def to_container(self, resolve: bool = True):
        def conv(v, where):
            if isinstance(v, dict):
                return {k: conv(x, f"{where}.{k}" if where else k) for k, x in v.items()}
            if isinstance(v, list):
                return [conv(x, where) for x in v]
            return self._resolve(v, where) if resolve else v

        return conv(self._data, self._path)

    def __repr__(self) -> str:
        return f"Cfg({self.to_container(resolve=False)!r})"


def compose(config_name: str = "basic", overrides: Sequence[str] = (), config_dir: Optional[str] = None) -> Cfg:
    """What ``hydra.compose`` / ``@hydra.main`` hand to ``main(cfg)``."""
    config_dir = config_dir or CONFIG_DIR
    root = _compose_file(config_dir, config_name)
    apply_overrides(root, list(overrides))
    return Cfg(root)


def parse_cli(argv: Sequence[str]) -> Tuple[str, Optional[str], List[str]]:
    """Splits a Hydra-style command line into (config_name, config_dir or None, overrides)."""
    name, cdir, ov = "basic", None, []
    it = iter(argv)
    for a in it:
        if a.startswith("--config-name="):
            name = a.split("=", 1)[1]
        elif a in ("--config-name", "-cn"):
            name = next(it)
        elif a.startswith("--config-path="):
            cdir = a.split("=", 1)[1]
        elif a in ("--config-path", "-cp"):
            cdir = next(it)
        else:
            ov.append(a)
    return name, cdir, ov


def workspace_kwargs(cfg: Cfg) -> Dict[str, Any]:
    """The ``**cfg`` splat of train.py:56-61 / infer.py:231-236 as plain containers (interpolations resolved; keys whose
    mandatory value was never given -- e.g. ``exp_name`` when composing programmatically -- are left out)."""
    out = {}
    for k in cfg.keys():
        try:
            v = cfg[k]
        except MissingMandatoryValue:
            continue
        out[k] = v.to_container() if isinstance(v, Cfg) else v
    return out
