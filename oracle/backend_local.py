"""Oracle restatement of the local backend, backend/local/backend.go:14-132
(TEST INFRASTRUCTURE ONLY): plain os calls around the C oracle's State."""
import os
import shutil

from . import oracle as O
from .records import render


class LocalBackend:
    def __init__(self, root_dir=None):
        self.root = root_dir or os.path.join(os.environ["HOME"], ".triton-kubernetes")   # homedir.Expand("~/.triton-kubernetes")
        os.makedirs(self.root, exist_ok=True)                                             # :35 os.MkdirAll

    def State(self, name):
        path = os.path.join(self.root, name, "main.tf.json")
        try:
            os.stat(path)
        except FileNotFoundError:                                                         # os.IsNotExist: ENOENT only
            return O.New(name, b"{}")                                                     # :51-55
        return O.New(name, open(path, "rb").read())                                      # :60-65

    def DeleteState(self, name):
        shutil.rmtree(os.path.join(self.root, name), ignore_errors=True)                 # :68-77 os.RemoveAll

    def PersistState(self, state):
        d = os.path.join(self.root, state.Name)
        os.makedirs(d, exist_ok=True)
        with open(os.path.join(d, "main.tf.json"), "wb") as f:                           # :94 ioutil.WriteFile(..., 0644)
            f.write(state.Bytes())

    def States(self):
        return sorted(e for e in os.listdir(self.root) if os.path.isdir(os.path.join(self.root, e)))   # :102-121

    def StateTerraformConfig(self, name):
        return "terraform.backend.local", render("localTerraformBackendConfig",
                                                 {"path": os.path.join(self.root, name, "terraform.tfstate")})   # :123-132
