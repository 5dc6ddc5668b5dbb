"""Host-side utilities: the reference's text protocol, validation and timers."""
from .protocol import (format_float, format_point, result_line, specify_problem_argv, specify_problem_stdin,
                       validate_input)
from . import trace
from .timing import CudaTimer, PhaseTimer

__all__ = ["format_float", "format_point", "result_line", "specify_problem_argv", "specify_problem_stdin",
           "validate_input", "CudaTimer", "PhaseTimer", "trace"]
