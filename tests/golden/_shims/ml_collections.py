"""Fixture: the sliver of ml_collections that config/sample_config.py:4-87 touches."""


class ConfigDict(dict):
    def __init__(self, initial_dictionary=None):
        super().__init__(initial_dictionary or {})

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError:
            raise AttributeError(k)

    def __setattr__(self, k, v):
        self[k] = v

    def get_ref(self, k):
        return self[k]
