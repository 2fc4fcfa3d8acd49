"""Checkpoint conventions of the reference's models/base_model.py:9-89: `<checkpoints_dir>/<name>/<epoch>_net_<label>.pth`
holding a plain state_dict whose keys and logical shapes equal the reference's, so files move both ways."""
import os

import torch


class BaseModel(torch.nn.Module):
    def name(self):
        return 'BaseModel'

    def initialize(self, opt):
        self.opt = opt
        self.gpu_ids = opt.gpu_ids
        self.isTrain = opt.isTrain
        self.save_dir = os.path.join(opt.checkpoints_dir, opt.name)
        # base_model.py:15 -- there is no CPU path here: a model built with gpu_ids=[] can be inspected
        # (state_dict, loss_names) but its forward raises
        self.device = ('cuda:%d' % self.gpu_ids[0]) if len(self.gpu_ids) > 0 else 'cpu'

    def save_network(self, network, network_label, epoch_label, gpu_ids):
        os.makedirs(self.save_dir, exist_ok=True)
        save_path = os.path.join(self.save_dir, '%s_net_%s.pth' % (epoch_label, network_label))
        torch.save({k: v.detach().cpu().contiguous() for k, v in network.state_dict().items()}, save_path)

    def load_network(self, network, network_label, epoch_label, save_dir=''):
        """strict load, else the subset of matching keys, else the keys whose sizes match (base_model.py:51-89)"""
        from . import networks
        save_path = os.path.join(save_dir or self.save_dir, '%s_net_%s.pth' % (epoch_label, network_label))
        if not os.path.isfile(save_path):
            print('%s not exists yet!' % save_path)
            if network_label == 'G':
                raise FileNotFoundError('Generator must exist!')
            return
        pretrained = torch.load(save_path, map_location='cpu')
        model_dict = network.state_dict()
        try:
            network.load_state_dict(pretrained)
        except RuntimeError:
            usable = {k: v for k, v in pretrained.items() if k in model_dict and v.size() == model_dict[k].size()}
            missing = sorted({k.split('.')[0] for k in model_dict if k not in usable})
            if missing:
                print('Pretrained network %s has fewer layers; The following are not initialized:' % network_label)
                print(missing)
            network.load_state_dict(usable, strict=False)
        networks.sync_shadows(network)
