"""Drop-in for the reference's upscaler.model (upscaling/upscaler/model.py) on the SRGAN hot path.

Same builder / loss / compile-helper names and argument meaning; the returned objects honour the Keras Model surface
the reference uses (predict, train_on_batch, save, input_shape, output_shape, layers, name, trainable).  Underneath,
every layer runs as hand-written sm_100a CUDA through libb2srgan.so — there is no Keras, TensorFlow or CPU path.
"""
import math
from abc import ABCMeta, abstractmethod

from .keras_compat import Adam, Input, Model  # noqa: F401  (re-exported: scripts import these from keras)
from . import networks
from .hdf5_io import load_model  # noqa: F401  (keras.models.load_model for the hot path's model families)


def make_upscaler_orig(output_image_shape, kernel_size=5, filters=64, upscale_factor=4, res_block_num=16):
    """model.py:267-295 — SRGAN generator; returns a Model whose input is output_image_shape // upscale_factor."""
    input_image_shape = (output_image_shape[0] // upscale_factor, output_image_shape[1] // upscale_factor,
                         output_image_shape[2])
    if output_image_shape[2] != 3:
        raise ValueError("make_upscaler_orig: the final conv emits 3 channels (model.py:290)")
    int(math.log(upscale_factor, 2))
    net = networks.UpscalerOrigNet(kernel_size=kernel_size, filters=filters, upscale_factor=upscale_factor,
                                   res_block_num=res_block_num)
    return Model(net=net, input_shape=input_image_shape, name="model_1")


def make_upscaler_skip_con(output_image_shape, kernel_size=5, filters=64, upscale_factor=4):
    """model.py:332-363 — skip-connection upscaler: orig trunk (16 residual blocks), 3x3 after-res conv, 224-filter 3x3
    Conv2DTranspose blocks, final conv over the concatenation of the bilinearly enlarged input and the features.
    Inference and training (compile_training_model / make_and_compile_gan*) run on the kernels of the `orig` path with the
    224-filter layers padded to 256 channels."""
    h, w, c = output_image_shape
    net = networks.UpscalerSkipConNet(kernel_size=kernel_size, filters=filters, upscale_factor=upscale_factor, res_block_num=16)
    return Model(net=net, input_shape=(h // upscale_factor, w // upscale_factor, c), name="model_1")


def _not_built(name, lines):
    raise NotImplementedError(
        "%s (reference upscaling/upscaler/model.py:%s) is not built on the B200 path yet; the generators that run are "
        "make_upscaler_orig, make_upscaler_skip_con and (inference) make_upscaler_attention / make_upscaler_incep_resnet "
        "(SURVEY.md 8f-3 lists the rest as next)" % (name, lines))


def make_upscaler_unetish(output_image_shape, kernel_size=5, upscale_factor=4, step_size=4, downscale_times=5,
                          initial_step_filter_count=32, dropout_rate=0.1):
    """model.py:563-636 (U-Net-ish generator with Dropout / Cropping2D / Concatenate).  Importable so that the reference
    scripts start (train.py:5, train_gan3.py:4); raises when called."""
    _not_built("make_upscaler_unetish", "563-636")


def make_upscaler_unetish_add(output_image_shape, kernel_size=5, upscale_factor=4, step_size=4, downscale_times=5,
                              initial_step_filter_count=48, dropout_rate=0.1):
    """model.py:642-722; raises when called."""
    _not_built("make_upscaler_unetish_add", "642-722")


def make_upscaler_unetish_complex(output_image_shape, kernel_size=5, upscale_factor=4, step_size=4, downscale_times=3,
                                  initial_step_filter_count=32, dropout_rate=0.1):
    """model.py:748-832; raises when called."""
    _not_built("make_upscaler_unetish_complex", "748-832")


def make_upscaler_attention(output_image_shape, kernel_size=5, filters=64, upscale_factor=4, res_block_num=16):
    """model.py:296-328 — the generator with sigmoid-gated residual blocks (residual_block_attention, :30-48) and gated
    upsampling blocks that add Conv2DTranspose(128, scale + 1, strides=scale)(atanh(0.99999 x)) of the network input
    (upsampling_block_attention, :78-98).  Inference (predict / save / load_model) runs on the B200 kernels; compiling it
    for training raises (its training step is not built)."""
    h, w, c = output_image_shape
    if c != 3:
        raise ValueError("make_upscaler_attention: the final conv emits 3 channels (model.py:322)")
    net = networks.UpscalerAttentionNet(kernel_size=kernel_size, filters=filters, upscale_factor=upscale_factor,
                                        res_block_num=res_block_num)
    return Model(net=net, input_shape=(h // upscale_factor, w // upscale_factor, c), name="model_1")


def make_upscaler_incep_resnet(output_image_shape, filters=64, upscale_factor=4, a_block_type='3path', a_block_num=5,
                               a_block_kernel=3, b_block_type='2path', b_block_num=10, b_block_kernel=7,
                               c_block_type='2path', c_block_num=5, c_block_kernel=3):
    """model.py:445-497 — generator of inception-style residual blocks (inception_resblock_3path / _2path, model.py:389-443:
    pre-activation BatchNorm -> PReLU -> Conv2D mini blocks with 1x1, kxk, 1xk, kx1 kernels).  Inference (predict / save /
    load_model) runs on the B200 kernels; compiling it for training raises (its training step is not built)."""
    h, w, c = output_image_shape
    if c != 3:
        raise ValueError("make_upscaler_incep_resnet: the final conv emits 3 channels (model.py:490)")
    if filters != 64:
        raise ValueError("the B200 path is built for filters=64 (reference default, model.py:446)")
    net = networks.UpscalerIncepResnetNet(filters=filters, upscale_factor=upscale_factor,
                                          a=(a_block_type, a_block_num, a_block_kernel), b=(b_block_type, b_block_num, b_block_kernel),
                                          c=(c_block_type, c_block_num, c_block_kernel))
    return Model(net=net, input_shape=(h // upscale_factor, w // upscale_factor, c), name="model_1")


def make_discriminator_simple_512(input_shape, activation='none'):
    """model.py:836-896 — 9 conv-BN-LeakyReLU(0.1) blocks (s1 then 8 x s2), Flatten, Dense 1024/32/1."""
    net = networks.DiscriminatorSimple512Net(tuple(input_shape), activation=activation)
    return Model(net=net, input_shape=tuple(input_shape), name="model_2")


def make_discriminator_thin_512(input_shape, activation='none'):
    """model.py:901-960 — make_discriminator_simple_512 with 128 filters in blocks 2..9."""
    net = networks.DiscriminatorThin512Net(tuple(input_shape), activation=activation)
    return Model(net=net, input_shape=tuple(input_shape), name="model_2")


def make_discriminator_sparse_512(input_shape, activation='none'):
    """model.py:964-1012 — six 5x5 padding='valid' conv-BN-LeakyReLU(0.1) blocks (strides 1,3,3,3,3,3), Flatten,
    Dense 128/32/1.  Inputs of at least 489 x 489."""
    net = networks.DiscriminatorSparse512Net(tuple(input_shape), activation=activation)
    return Model(net=net, input_shape=tuple(input_shape), name="model_2")


# ------------------------------------------------------------------------------------------------ content losses
class _VggLossBase(object):
    """Frozen VGG19 -> block5_conv4 features of y_true and y_pred (model.py:101-157).  `vgg19` may be a Model wrapping
    a Vgg19Net, a Vgg19Net, or a dict of Keras-layout weights; with None the network gets seeded He-normal weights —
    the ImageNet weights keras.applications downloads at run time (model.py:108) are unavailable offline."""
    kind = "vgg"

    def __init__(self, image_shape, rate=0.0, vgg19=None):
        self.image_shape = image_shape
        self.rate = rate
        if isinstance(vgg19, Model):
            net = vgg19.net
        elif isinstance(vgg19, networks.Vgg19Net):
            net = vgg19
        elif isinstance(vgg19, dict):
            net = networks.Vgg19Net(weights=vgg19)
        else:
            net = networks.Vgg19Net(weights=networks.init_vgg19_weights())
        self.net = net
        self.model = Model(net=net, input_shape=tuple(image_shape), name="vgg19_block5_conv4")
        self.model.trainable = False

    def loss(self, y_true, y_pred):
        """Evaluates the loss on two host batches (the symbolic use inside compile is resolved by the training engine)."""
        from . import training
        return training.evaluate_content_loss(self, y_true, y_pred)


class VGG_LOSS(_VggLossBase):
    """model.py:101-117: mean((V(t) - V(p))^2)."""
    kind = "vgg"

    def __init__(self, image_shape, vgg19=None):
        super().__init__(image_shape, 0.0, vgg19)


class VGG_MSE_LOSS(_VggLossBase):
    """model.py:120-137: mean((V(t) - V(p))^2) + rate * mean((t - p)^2)."""
    kind = "vgg-mse"

    def __init__(self, image_shape, mse_loss_rate=0.1, vgg19=None):
        super().__init__(image_shape, mse_loss_rate, vgg19)
        self.mse_loss_rate = mse_loss_rate


class VGG_MAE_LOSS(_VggLossBase):
    """model.py:140-157: mean(|V(t) - V(p)|) + rate * mean(|t - p|)."""
    kind = "vgg-mae"

    def __init__(self, image_shape, mae_loss_rate=0.1, vgg19=None):
        super().__init__(image_shape, mae_loss_rate, vgg19)
        self.mae_loss_rate = mae_loss_rate


# ------------------------------------------------------------------------------------------------ adversarial losses
def wasserstein_loss(y_true, y_pred):
    """model.py:159-160: mean(y_true * y_pred).  Inside compile the engine recognises this function by identity; called
    on host arrays it evaluates directly."""
    import numpy as np
    return float(np.mean(np.asarray(y_true, np.float64) * np.asarray(y_pred, np.float64)))


class GanLosses(metaclass=ABCMeta):
    """model.py:166-212."""
    family = None

    def __init__(self, loss_activation='log-sigm', real_output=None, fake_output=None):
        self._real_output = real_output
        self._fake_output = fake_output
        self.loss_activation_name = loss_activation if loss_activation in ('sigmoid', 'log-sigm', 'tanh', 'bi-log') else 'none'

    @property
    def real_output(self):
        return self._real_output

    @real_output.setter
    def real_output(self, real_output):
        self._real_output = real_output

    @property
    def fake_output(self):
        return self._fake_output

    @fake_output.setter
    def fake_output(self, fake_output):
        self._fake_output = fake_output

    @property
    @abstractmethod
    def discriminator_loss(self):
        pass

    @property
    @abstractmethod
    def generator_loss(self):
        pass


class WassersteinLosses(GanLosses):
    """model.py:215-235: D loss mean(real) - mean(fake); G loss mean(fake)."""
    family = "wasserstein"

    @property
    def discriminator_loss(self):
        from .training import GanLossFn
        return GanLossFn(self, "disc")

    @property
    def generator_loss(self):
        from .training import GanLossFn
        return GanLossFn(self, "gen")


class RelativisticLosses(GanLosses):
    """model.py:239-261: act(mean(real) - mean(fake)) / act(mean(fake) - mean(real))."""
    family = "relativistic"

    @property
    def discriminator_loss(self):
        from .training import GanLossFn
        return GanLossFn(self, "disc")

    @property
    def generator_loss(self):
        from .training import GanLossFn
        return GanLossFn(self, "gen")


# ------------------------------------------------------------------------------------------------ compile helpers
def make_and_compile_gan(generator, discriminator, input_shape, output_shape, content_loss, content_loss_weight,
                         discriminator_loss, discriminator_loss_weight, optimizer=Adam()):
    """model.py:1017-1051.  The default optimizer is one object created at import, shared by every compile call."""
    generator_input_layer = Input(shape=input_shape)
    generator_layer = generator(generator_input_layer)
    generator_training_model = Model(inputs=generator_input_layer, outputs=generator_layer)
    generator_training_model.compile(loss=content_loss, optimizer=optimizer)

    discriminator.trainable = True
    discriminator_input_layer = Input(shape=output_shape)
    discriminator_layer = discriminator(discriminator_input_layer)
    discriminator_training_model = Model(inputs=discriminator_input_layer, outputs=discriminator_layer)
    discriminator_training_model.compile(loss=discriminator_loss, optimizer=optimizer)

    discriminator.trainable = False
    gan_input_layer = Input(shape=input_shape)
    gan_generator_layer = generator(gan_input_layer)
    gan_discriminator_layer = discriminator(gan_generator_layer)
    gan_training_model = Model(inputs=gan_input_layer, outputs=[gan_generator_layer, gan_discriminator_layer])
    gan_training_model.compile(loss=[content_loss, discriminator_loss],
                               loss_weights=[content_loss_weight, discriminator_loss_weight], optimizer=optimizer)
    return generator_training_model, discriminator_training_model, gan_training_model


def make_and_compile_gan2(generator, discriminator, input_shape, output_shape, content_loss, content_loss_weight,
                          discriminator_losses, discriminator_loss_weight, optimizer=Adam()):
    """model.py:1057-1125 — two-input discriminator model, losses built from symbolic real / fake outputs."""
    generator_input = Input(shape=input_shape)
    generator_output = generator(generator_input)
    generator_training_model = Model(inputs=generator_input, outputs=generator_output)
    generator_training_model.compile(loss=content_loss, optimizer=optimizer)

    discriminator.trainable = True
    discriminator_real_input = Input(shape=output_shape)
    discriminator_real_output = discriminator(discriminator_real_input)
    discriminator_fake_input = Input(shape=output_shape)
    discriminator_fake_output = discriminator(discriminator_fake_input)
    disc_loss = discriminator_losses()
    disc_loss.real_output = discriminator_real_output
    disc_loss.fake_output = discriminator_fake_output
    disc_loss = disc_loss.discriminator_loss
    discriminator_training_model = Model(inputs=[discriminator_real_input, discriminator_fake_input],
                                         outputs=discriminator_fake_output)
    discriminator_training_model.compile(loss=disc_loss, optimizer=optimizer)

    discriminator.trainable = False
    gan_real_input = Input(shape=output_shape)
    gan_disc_real_output = discriminator(gan_real_input)
    gan_input_layer = Input(shape=input_shape)
    gan_generator_output = generator(gan_input_layer)
    gan_disc_fake_output = discriminator(gan_generator_output)
    gan_loss = discriminator_losses()
    gan_loss.real_output = gan_disc_real_output
    gan_loss.fake_output = gan_disc_fake_output
    gan_loss = gan_loss.generator_loss
    gan_training_model = Model(inputs=[gan_input_layer, gan_real_input], outputs=[gan_generator_output, gan_disc_fake_output])
    gan_training_model.compile(loss=[content_loss, gan_loss], loss_weights=[content_loss_weight, discriminator_loss_weight],
                               optimizer=optimizer)
    return generator_training_model, discriminator_training_model, gan_training_model


def compile_training_model(upscaler, loss, optimizer=Adam()):
    """model.py:1130-1137."""
    input_layer = Input(shape=upscaler.input_shape[1:4])
    upscaled_layer = upscaler(input_layer)
    upscaler_training_model = Model(inputs=input_layer, outputs=upscaled_layer)
    upscaler_training_model.compile(loss=loss, optimizer=optimizer)
    return upscaler_training_model
