"""Plumbing workload: print (and save) the rank's PADDLE_* environment, then
exit.  Used by the CPU process tests and examples/env_dump.yaml to show that a
rank sees exactly the variables the reference injects
(controllers/paddlejob_helper.go:253-277,307-347)."""
import json
import os
import sys
import time


def main() -> int:
    keys = sorted(k for k in os.environ if k.startswith(("PADDLE_", "TRAINING_ROLE", "TRAINER_PORTS_NUM", "POD_IP",
                                                         "PDJ_", "CUDA_VISIBLE_DEVICES")))
    env = {k: os.environ[k] for k in keys}
    out = os.environ.get("PDJ_ENV_DUMP_DIR")
    if out:
        os.makedirs(out, exist_ok=True)
        with open(os.path.join(out, os.environ.get("PDJ_POD_NAME", "rank") + ".json"), "w") as f:
            json.dump(env, f, indent=1, sort_keys=True)
    print(json.dumps(env, sort_keys=True), flush=True)
    time.sleep(float(os.environ.get("PDJ_ENV_DUMP_SLEEP", "0.2")))
    return int(os.environ.get("PDJ_ENV_DUMP_EXIT", "0"))


if __name__ == "__main__":
    sys.exit(main())
