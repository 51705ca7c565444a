"""`samgraph` as the training scripts import it -- `import xgnn_b200.samgraph.torch as sam` in place of
`import samgraph.torch as sam` (example/samgraph/sgnn/train_graphsage.py:25).

common.py mirrors samgraph/common/__init__.py (enums, log-item numbering, the ctypes `_basics` wrapper),
torch.py mirrors samgraph/torch/adapter.py.  Both bind libxgnn_b200.so's `samgraph_*` C-ABI
(include/samgraph_abi.h); there is no other backend."""
from .common import *  # noqa: F401,F403
