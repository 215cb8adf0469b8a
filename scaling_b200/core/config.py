"""Config base class (reference: core/config/base.py:26-152).

Frozen, ``extra="forbid"`` pydantic models; YAML (JSON subset + comments) or JSON in, JSON out; recursive dict
overwrite; and a self-documenting YAML template generated from the field descriptions.
"""
from __future__ import annotations

import json
from enum import Enum
from pathlib import Path
from typing import Any, Optional, Type, TypeVar, Union

import yaml
from pydantic import BaseModel, ConfigDict

T = TypeVar("T", bound="BaseConfig")


def overwrite_recursive(d: dict, d_new: dict) -> None:
    """Merge ``d_new`` into ``d`` in place; nested dicts are merged key by key, everything else is replaced."""
    for key, value in d_new.items():
        if isinstance(value, dict):
            if not isinstance(d.get(key), dict):
                d[key] = {}
            overwrite_recursive(d[key], value)
        else:
            d[key] = value


def _jsonable(x: Any) -> Any:
    if isinstance(x, dict):
        return {k: _jsonable(v) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return [_jsonable(v) for v in x]
    if isinstance(x, Path):
        return str(x)
    if isinstance(x, Enum):
        return x.value
    return x


class BaseConfig(BaseModel):
    """Base config: immutable, unknown keys rejected, JSON/YAML round trip."""

    model_config = ConfigDict(extra="forbid", frozen=True, protected_namespaces=())

    def as_dict(self) -> dict[str, Any]:
        return _jsonable(self.model_dump())

    def as_str(self) -> str:
        return json.dumps(self.as_dict())

    @classmethod
    def from_dict(cls: Type[T], d: dict, overwrite_values: Optional[dict] = None) -> T:
        if overwrite_values is not None:
            overwrite_recursive(d, overwrite_values)
        return cls(**d)

    @classmethod
    def from_str(cls: Type[T], s: str) -> T:
        return cls.from_dict(json.loads(s))

    @classmethod
    def from_yaml(cls: Type[T], yml_filename: Union[str, Path], overwrite_values: Optional[dict] = None) -> T:
        with open(yml_filename, encoding="utf-8") as f:
            d = yaml.load(f, Loader=yaml.FullLoader)
        if overwrite_values is not None:
            overwrite_recursive(d, overwrite_values)
        return cls.from_dict(d)

    def save(self, out_file: Union[str, Path], indent: int = 4) -> None:
        with open(out_file, "w", encoding="utf-8") as f:
            json.dump(self.as_dict(), f, indent=indent)

    # ---- template --------------------------------------------------------------------------------
    @classmethod
    def get_template_str(cls, indent: int = 4, level: int = 1) -> str:
        pad = " " * indent * level

        def comment(text: Optional[str]) -> str:
            if not text or not text.strip():
                return ""
            return "".join(f"{pad}# {line}\n" for line in text.split("\n"))

        names = list(cls.model_json_schema(by_alias=False)["properties"].keys())
        out = "{\n" + comment(cls.__name__) + comment(cls.__doc__)
        for i, name in enumerate(names):
            info = cls.model_fields[name]
            out += pad + "\n" + comment(info.description) + pad + f'"{name}": '
            default = info.default
            if isinstance(default, BaseConfig):
                out += type(default).get_template_str(indent=indent, level=level + 1)
            elif isinstance(default, Enum):
                out += json.dumps(default.value)
            elif default is None or type(default).__name__ == "PydanticUndefinedType":
                out += "null"  # required fields have no default: the template shows where to fill them in
            else:
                try:
                    out += json.dumps(_jsonable(default))
                except TypeError:
                    out += "null"
            if i != len(names) - 1:
                out += ","
            out += "\n"
        return out + " " * indent * (level - 1) + "}"

    def as_annotated_str(self, indent: int = 4, level: int = 1) -> str:
        """This instance's values rendered like the template: every field preceded by its description as a comment.
        The result is YAML (JSON with ``#`` comments) that ``from_yaml`` reads back."""
        pad = " " * indent * level
        names = list(type(self).model_fields.keys())
        out = "{\n" + f"{pad}# {type(self).__name__}\n"
        for i, name in enumerate(names):
            info = type(self).model_fields[name]
            value = getattr(self, name)
            for line in (info.description or "").split("\n"):
                if line.strip():
                    out += f"{pad}# {line}\n"
            out += pad + f'"{name}": '
            if isinstance(value, BaseConfig):
                out += value.as_annotated_str(indent=indent, level=level + 1)
            else:
                out += json.dumps(_jsonable(value))
            out += ("," if i != len(names) - 1 else "") + "\n\n"
        return out.rstrip("\n") + "\n" + " " * indent * (level - 1) + "}"

    @classmethod
    def save_template(cls, out_file: Union[str, Path], indent: int = 4) -> None:
        with open(out_file, "w", encoding="utf-8") as f:
            f.write(cls.get_template_str(indent=indent))
