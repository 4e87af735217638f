"""Controller base class (mirror of vehicle_control/controllers/controller.py:6-19)."""
from abc import ABC, abstractmethod


class Controller(ABC):
    @abstractmethod
    def command(self, *args, **kwargs):
        """Compute the next action."""
