"""pmtech_b200 — B200-native replacement for pmtech's per-frame ECS scene-update hot path
(put::ecs::update_scene loops A-C + per-view filter / frustum cull / draw-list build).

  csrc/      CUDA kernels (sm_100a) and the C ABI of include/pmtech_b200.h
  capi.py    ctypes binding of the C ABI
  scene.py   DeviceScene: RAII handle over the device mirrors
  ecs.py     put::ecs-shaped host API (update_scene, filter_entities_scalar, frustum_cull_*)
  synth.py   synthetic scenes / views of the BASELINE.json configs
  shard.py   root-subtree sharding over several GPUs (torch.distributed plumbing)

There is no CPU implementation of the path in this package: without the built library or
without a CUDA device every entry point raises."""
__version__ = "0.1"
