"""The settings store the flows read: spf13/viper as the reference uses it (cmd/root.go:48-67).

Precedence (viper.find): values set in-process (viper.Set) > environment (AutomaticEnv: the key in
upper case) > the --config YAML file.  Keys are case-insensitive.  A value set to None does not
count as set and does not shadow the sources below it (create/cluster.go:177-208 relies on this
when it fans a node entry into the global store).
"""
from __future__ import annotations

import os


class Config:
    def __init__(self, file_values=None, environ=None):
        self._file = {str(k).lower(): v for k, v in (file_values or {}).items()}
        self._override = {}
        self._env = os.environ if environ is None else environ

    @classmethod
    def from_yaml(cls, path, environ=None):
        import yaml
        with open(path) as f:
            data = yaml.safe_load(f) or {}
        if not isinstance(data, dict):
            raise ValueError("config file must be a YAML mapping")
        return cls(data, environ)

    def Set(self, key, value):
        self._override[key.lower()] = value

    def Reset(self):
        self._override.clear()

    def Get(self, key):
        k = key.lower()
        v = self._override.get(k)
        if v is not None:
            return v
        e = self._env.get(k.upper())
        if e is not None:
            return e
        return self._file.get(k)

    def IsSet(self, key) -> bool:
        return self.Get(key) is not None

    def GetString(self, key) -> str:
        v = self.Get(key)
        if v is None:
            return ""
        if isinstance(v, bool):
            return "true" if v else "false"
        if isinstance(v, (list, dict)):
            return ""
        return str(v)

    def GetBool(self, key) -> bool:
        v = self.Get(key)
        if isinstance(v, str):
            return v.lower() in ("1", "t", "true")
        return bool(v)

    def GetStringSlice(self, key) -> list:
        v = self.Get(key)
        if v is None:
            return []
        if isinstance(v, str):
            return v.split()
        return [str(x) for x in v]
