"""`python -m triton-kubernetes_b200 [--config FILE] --non-interactive create|destroy|get manager|cluster|node`

The reference's command line (cmd/root.go:39-67, cmd/create.go:15-67, cmd/destroy.go:14-75,
cmd/get.go:14-66) as a thin driver over flows/, for the non-interactive runs BASELINE.json's configs
#2 and #3 name.  What it keeps of the reference: the two persistent flags, the settings precedence
(viper: in-process > environment > --config YAML), the argument checks with their error texts, the
lines the commands print, and exit status 1 with the flow's error text on stdout.  What it leaves
out: interactive prompts (a run without --non-interactive is refused), cloud API providers and the
Manta backend (the flows refuse them by name).  --store points the local backend somewhere other
than ~/.triton-kubernetes (tests use it); the reference has no such flag.
"""
from __future__ import annotations

import os
import sys

VALID = {"create": ("manager", "cluster", "node"), "destroy": ("manager", "cluster", "node"), "get": ("manager", "cluster")}
SPELLED = {"create": "create", "destroy": "destory", "get": "get"}     # cmd/destroy.go:22,31 spells it that way


def main(argv=None, out=print) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    cfg_file, store, non_interactive, words = "", None, False, []
    i = 0
    while i < len(argv):
        a = argv[i]
        if a == "--non-interactive":
            non_interactive = True
        elif a.startswith("--non-interactive="):
            non_interactive = a.split("=", 1)[1].lower() in ("1", "t", "true")
        elif a in ("--config", "--store"):
            if i + 1 >= len(argv):
                out("Error: flag needs an argument: %s" % a)
                return 1
            i += 1
            if a == "--config":
                cfg_file = argv[i]
            else:
                store = argv[i]
        elif a.startswith("--config="):
            cfg_file = a.split("=", 1)[1]
        elif a.startswith("--store="):
            store = a.split("=", 1)[1]
        elif a.startswith("-"):
            out("Error: unknown flag: %s" % a)
            return 1
        else:
            words.append(a)
        i += 1
    if not words or words[0] not in VALID:
        out("usage: python -m triton-kubernetes_b200 [--config FILE] --non-interactive create|destroy|get manager|cluster|node")
        return 1
    verb, args = words[0], words[1:]
    if len(args) != 1:
        out('Error: "triton-kubernetes %s" requires one argument' % SPELLED[verb])
        return 1
    if args[0] not in VALID[verb]:
        out('Error: invalid argument "%s" for "triton-kubernetes %s"' % (args[0], SPELLED[verb]))
        return 1

    from . import LocalBackend, flows, records
    from . import state as tk_state
    import importlib
    pkg = importlib.import_module(__package__)

    # cmd/root.go:48-67
    if non_interactive:
        out("Running in non interactive mode")
    default_cfg = os.path.join(os.path.expanduser("~"), ".triton-kubernetes.yaml")
    path = cfg_file or (default_cfg if os.path.exists(default_cfg) else "")
    config = flows.Config()
    if path:
        try:
            config = flows.Config.from_yaml(path)
            out("Using config file: %s" % path)
        except (OSError, ValueError) as e:
            if cfg_file:          # viper.ReadInConfig's error is dropped by the reference; an explicit --config that cannot be read is worth a word
                out("Cannot read config file %s: %s" % (cfg_file, e))
    config.Set("non-interactive", non_interactive)
    if not non_interactive:
        out("interactive prompts are not part of this build: run with --non-interactive and a --config file")
        return 1

    env = flows.Env(state=pkg, backend=None, render=records.render, module_source=records.module_source, config=config)
    try:
        env.backend = flows.backend.prompt_for_backend(config, lambda: LocalBackend(store))       # util.PromptForBackend
        out("%s %s called" % (verb, args[0]))
        fn = {("create", "manager"): flows.create.new_manager, ("create", "cluster"): flows.create.new_cluster,
              ("create", "node"): flows.create.new_node, ("destroy", "manager"): flows.destroy.delete_manager,
              ("destroy", "cluster"): flows.destroy.delete_cluster, ("destroy", "node"): flows.destroy.delete_node,
              ("get", "manager"): flows.get.get_manager, ("get", "cluster"): flows.get.get_cluster}[(verb, args[0])]
        fn(env)
    except (flows.FlowError, tk_state.StateError) as e:
        for line in env.out:
            out(line)
        out(str(e))
        return 1
    for line in env.out:
        out(line)
    return 0


if __name__ == "__main__":
    sys.exit(main())
