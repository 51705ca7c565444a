"""arch5: decoupled sampler / trainer processes through the samgraph_* C-ABI.  Kept in its own module, sorted
after the single-process and arch6 flows: it forks 2-4 CUDA processes per case."""
import pytest

from test_samgraph_api import _drive, dataset  # noqa: F401  (module-scoped dataset fixture)


@pytest.mark.gpu
@pytest.mark.parametrize("sample_type,samplers,trainers", [(0, 1, 1), (0, 2, 2), (7, 2, 1), (3, 1, 2)])
def test_arch5_decoupled_sampler_and_trainer_processes(dataset, sample_type, samplers, trainers):
    """arch5: samplers and trainers are separate processes (separate GPUs when the box has them); a Task's
    topology crosses as ONE peer copy out of the sampler's HBM, only a descriptor goes through host shared memory
    (reference: D2H + 50 MB host ring + H2D, memory_queue.cc / task_queue.cc).  Every Task reaches exactly one trainer
    and carries the oracle's blocks and CPUExtract's rows bit for bit."""
    import torch
    _drive("arch5", dataset, sample_type, samplers, trainers, env={"XG_TEST_NUM_GPU": str(max(1, torch.cuda.device_count())), "SAMGRAPH_QUEUE_TIMEOUT": "60"})
