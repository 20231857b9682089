"""The plug-in interface a circuit back-end implements.

Same abstract methods, names and call protocol as the reference's
``CircuitBuilder`` (``quompiler/circuits/qbuilder.py:9-49``): ``register`` may be
called repeatedly, ``build_gate`` is called in circuit time order, ``build_group``
is an optional hook for inner tree nodes, ``finish`` may be called more than once.
"""
from abc import ABC, abstractmethod


class CircuitBuilder(ABC):
    @abstractmethod
    def register(self, qspace) -> None:
        ...

    @abstractmethod
    def build_gate(self, m) -> None:
        ...

    def build_group(self, m) -> None:
        pass

    @abstractmethod
    def finish(self, optimized=False) -> object:
        ...

    @abstractmethod
    def all_qubits(self) -> list:
        ...
