"""Built-in example functions (reference: ml/experiments/kubeml/function_*.py).  Each module has
a `main()` that builds network + dataset + KubeModel and returns `kubenet.start()`, exactly the
shape of a user-written function (README.md:139-246 of the reference)."""
