"""Mirror of clients/preprocess/base_preprocess.py:4-10."""
from abc import ABC, abstractmethod


class Preprocess(ABC):
    @abstractmethod
    def preprocess(self):
        """ """
