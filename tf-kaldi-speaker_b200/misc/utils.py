"""Config objects with the reference's interface (misc/utils.py:13-61): JSON -> attributes, `.dict` view.
Layer code injects defaults lazily by mutating `params.dict` (model/tdnn.py:114-116,152-154,163-164,173-174);
that behaviour is part of the observable API and is kept."""
import json


class Params():
    """Loads hyperparameters from a json file (reference: misc/utils.py:13-41)."""

    def __init__(self, json_path):
        self.update(json_path)

    def save(self, json_path):
        with open(json_path, 'w') as f:
            json.dump(self.__dict__, f, indent=4)

    def update(self, json_path):
        with open(json_path) as f:
            params = json.load(f)
            self.__dict__.update(params)

    @property
    def dict(self):
        return self.__dict__


class ParamsPlain():
    """Hyperparameters set by hand (reference: misc/utils.py:44-61)."""

    def __init__(self):
        pass

    @property
    def dict(self):
        return self.__dict__
