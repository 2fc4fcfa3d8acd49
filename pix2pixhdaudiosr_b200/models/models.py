"""models/models.py:3-20 of the reference: `create_model(opt)`.  The DataParallel wrap of :17-18 is replaced by one
process per GPU (pix2pixhdaudiosr_b200/train.py), so the model is returned unwrapped."""


def create_model(opt):
    if opt.model != 'pix2pixHD':
        raise NotImplementedError('model [%s]: only pix2pixHD is on the hot path (ui_model is image-only legacy)' % opt.model)
    from .pix2pixHD_model import InferenceModel, Pix2PixHDModel
    model = Pix2PixHDModel() if opt.isTrain else InferenceModel()
    model.initialize(opt)
    if opt.verbose:
        print("model [%s] was created" % (model.name()))
    return model
