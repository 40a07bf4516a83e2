"""shell/run_shell_cmd.go:8-29 and shell/run_terraform.go:11-80: terraform is exec'ed in a
temporary directory that holds State.Bytes() as main.tf.json."""
from __future__ import annotations

import os
import shutil
import subprocess
import tempfile

from .env import FlowError


def run_shell_command(env, working_dir, name, *args):
    try:
        rc = subprocess.call([name, *args], cwd=working_dir)
    except OSError as e:
        raise FlowError('exec: "%s": %s' % (name, e.strerror or e))
    if rc != 0:
        raise FlowError("exit status %d" % rc)


def _with_state(env, state, *commands):
    temp = tempfile.mkdtemp(prefix="triton-kubernetes-")           # ioutil.TempDir("", "triton-kubernetes-")
    try:
        with open(os.path.join(temp, "main.tf.json"), "wb") as f:   # run_terraform.go:21, :56
            f.write(state.Bytes())
        os.chmod(os.path.join(temp, "main.tf.json"), 0o644)
        for cmd in commands:
            run_shell_command(env, temp, env.terraform, *cmd)
    finally:
        shutil.rmtree(temp, ignore_errors=True)


def run_terraform_apply_with_state(env, state):
    _with_state(env, state, ["init", "-force-copy"], ["apply", "-auto-approve"])


def run_terraform_destroy_with_state(env, state, args):
    _with_state(env, state, ["init", "-force-copy"], ["destroy", "-force", *args])


def run_terraform_output_with_state(env, state, module_key):
    """get/cluster.go:109-138, get/manager.go:63-93."""
    _with_state(env, state, ["init", "-force-copy"], ["output", "-module", module_key])
