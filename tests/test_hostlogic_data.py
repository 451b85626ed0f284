"""CPU: the PyG-free batch container and loader (deltaconv_b200.utils.Batch / DataLoader)."""
import torch


def test_batch_from_data_list_and_loader():
    from deltaconv_b200.utils import Data, Batch, DataLoader
    torch.manual_seed(0)
    clouds = [Data(pos=torch.randn(n, 3), norm=torch.randn(n, 3), y=torch.tensor([i % 3])) for i, n in enumerate((5, 9, 1, 7))]
    b = Batch.from_data_list(clouds)
    assert b.pos.shape == (22, 3) and b.norm.shape == (22, 3) and b.y.tolist() == [0, 1, 2, 0]
    assert b.batch.tolist() == [0] * 5 + [1] * 9 + [2] + [3] * 7
    assert (b._num_clouds, b._max_points, b.num_graphs) == (4, 9, 4)
    assert torch.equal(b.pos[5:14], clouds[1].pos)
    seg = Batch.from_data_list([Data(pos=torch.randn(n, 3), y=torch.randint(0, 4, (n,)), category=torch.tensor([i])) for i, n in enumerate((6, 4))])
    assert seg.y.shape == (10,) and seg.category.tolist() == [0, 1]
    moved = b.to('cpu')
    assert isinstance(moved, Batch) and moved._num_clouds == 4
    loader = DataLoader(clouds, batch_size=2)
    batches = list(loader)
    assert len(batches) == 2 and batches[0].pos.shape == (14, 3) and batches[1].batch.tolist() == [0] + [1] * 7
