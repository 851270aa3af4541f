"""paddle_operator_b200 - a Blackwell-native, single-box PaddleJob runtime.

Same public API as Paddle Operator (the `PaddleJob` batch.paddlepaddle.org/v1
documents, the PADDLE_* rank environment, status phases), re-targeted from
Kubernetes pods to rank processes on one 8xB200 machine, plus the compute and
communication plane the reference leaves to the user image.

Sub-packages
  api         PaddleJob types + YAML loader           (ref: api/v1/*.go)
  controller  helpers, object store, reconciler        (ref: controllers/*.go)
  runtime     supervisor, CLI/manager, etcd store      (ref: main.go, kubelet, etcd)
  ops         hand-written sm_100a kernels (ctypes)    (no counterpart in the reference)
  parallel    symmetric memory, collective / PS engines
  models      ResNet-50, Wide&Deep
  workloads   rank programs and benchmarks
"""
__version__ = "0.1.0"
