"""TEST INFRASTRUCTURE ONLY: the CPU oracle and the bindings to the compiled reference. Never imported by svtrek_b200."""
