"""`utils.dataset` of the reference: the PNG-pair dataset with the same discovery rules (utils/dataset.py:13-136), as a
``torch.utils.data.Dataset`` of (low, full) float32 tensors so that the reference's DataLoader code keeps working."""
import torch
from torch.utils.data import Dataset

from mri_superresolution_b200.data import PairedPNGDataset


class MRISuperResDataset(Dataset):
    def __init__(self, full_res_dir, low_res_dir, transform=None, augmentation=True, augmentation_params=None, normalize=True,
                 cache_size=100):
        self._ds = PairedPNGDataset(full_res_dir, low_res_dir, cache=cache_size > 0)
        self.full_res_files, self.valid_pairs = self._ds.full_res_files, self._ds.valid_pairs
        self.subjects, self.metadata = self._ds.subjects, self._ds.metadata
        self.augmentation = augmentation            # applied on the device by drivers.train(cfg.augmentation), not per item

    def __len__(self):
        return len(self._ds)

    def __getitem__(self, idx):
        low, full = self._ds.load(idx)
        return (torch.from_numpy(low).float().div(255)[None], torch.from_numpy(full).float().div(255)[None])

    def get_subject_indices(self, subject_id):
        return self._ds.get_subject_indices(subject_id)
