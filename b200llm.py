"""Import alias: ``import b200llm`` -> the package directory whose (mandated) name contains
hyphens and therefore cannot appear in an ``import`` statement. Submodules are aliased by the
package's own __init__ (``b200llm.ops`` is the same module object as ``<real>.ops``)."""
import importlib
import sys

REAL_NAME = "distributed-real-time-chat-and-collaboration-tool_b200"
_pkg = importlib.import_module(REAL_NAME)
sys.modules[__name__] = _pkg
