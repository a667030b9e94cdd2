"""Import shim: the package directory is named `flink-jpmml_b200/` (not a valid Python identifier), so this module
presents it as the importable package `flink_jpmml_b200` (sub-modules: capi, api, pmmlgen, ...)."""
import os as _os

__path__ = [_os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "flink-jpmml_b200")]
with open(_os.path.join(__path__[0], "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(__path__[0], "__init__.py"), "exec"))
