"""Graph base: inputs, embedding, predict/accuracy, optimizer, train/eval loops - host mirror of the
reference's ``model/base.py``.

TF's "build the graph once, then ``sess.run`` it" becomes "build the pipeline once, then step it":
``build()`` allocates the parameter store and the static buffers of an ``EpisodePipeline``;
``feed(inputs, query_label, keep_prob)`` plays the role of ``feed_dict``; ``forward`` /
``build_predict`` / ``build_accuracy`` / ``build_loss`` / ``build_optimize`` then *run* their part of
the step on the fed batch and refresh the attributes the reference exposes (``probs``, ``alphas``,
``predict_test``, ``predict``, ``accuracy``, ``loss``, ``learning_rate``, ``global_step``).
``train`` / ``eval`` keep the reference's signatures, print formats and best-checkpoint policy.
"""
from __future__ import annotations

from os import makedirs
from os.path import exists, join

import numpy as np
import torch

from ..engine import POS_DIM, EpisodeConfig, EpisodePipeline, ParamStore, init_tf_params, tf_param_shapes


class Base:
    def __init__(self, kwds):
        # reference keyword set (model/base.py:13-26)
        self.sequence_length = kwds.get("sequence_length", 40)
        self.vocab_size = kwds.get("vocab_size", None)
        self.embed_size = kwds.get("embed_size", 50)
        self.hidden_size = kwds.get("hidden_size", 100)
        self.is_training = kwds.get("is_training", True)
        self.learning_rate0 = kwds.get("learning_rate", 0.001)
        self.learning_rate = self.learning_rate0
        self.initializer = kwds.get("initializer", None)     # accepted for signature compatibility
        self.decay_steps = kwds.get("decay_steps", 100)
        self.decay_rate = kwds.get("decay_rate", 0.9)
        self.l2_lambda = kwds.get("l2_lambda", 0.00001)      # dead in the reference too (graph.py:108-109)
        self.embed = kwds.get("pred_embed", None)
        self.pos_embedding_dim = POS_DIM
        self.keepProb = kwds.get("keep_prob", 0.9)
        self.alphas = None
        # extensions (not in the reference): episodes per step, NTN slices, precision mode, seeds
        self.batch_size = kwds.get("batch_size", 1)
        self.out_size = kwds.get("out_size", 100)
        self.mode = kwds.get("mode", "fp32")
        self.seed = kwds.get("seed", 0)
        # cuda_graph: run train() through EpisodeTrainer (device sampler in the loop, one CUDA graph per step,
        # loss / accuracy read every 100 steps); implied when launched under torchrun with WORLD_SIZE > 1
        self.use_cuda_graph = kwds.get("cuda_graph", False)
        self.scalar_log = kwds.get("scalar_log", None)       # JSONL file for the tf.summary scalars (base.py:110-113)

    # -------------------------------------------------------------------------------- build
    def build(self):
        self.initial_params()
        self.global_step = 0
        self.probs = self.loss = self.accuracy = self.predict = self.predict_test = None
        self.summary = {}

    def initial_params(self):
        """Placeholders -> static device buffers; embedding tables and all other trainables -> one
        flat parameter store (reference model/base.py:38-74)."""
        if self.embed is not None:
            emb = torch.as_tensor(np.asarray(self.embed), dtype=torch.float32)
            print("use word embedding, size:", tuple(emb.shape))
            self.embed_size = int(emb.shape[1])
            vocab_rows = None
        else:
            if self.vocab_size is None:
                raise ValueError("vocab_size is required when no pred_embed is given")
            emb, vocab_rows = None, int(self.vocab_size)
        self.cfg = EpisodeConfig(N=self.num_classes, K=self.support_num_per_class, Q=self.query_num_per_class,
                                 B=self.batch_size, L=self.sequence_length, E=self.embed_size,
                                 u=self.hidden_size, S=self.out_size, keep_prob=float(self.keepProb),
                                 mode=self.mode)
        shapes = tf_param_shapes(self.cfg.N, self.cfg.L, self.cfg.E, self.cfg.u, self.cfg.S, vocab_rows)
        self.store = ParamStore(shapes)
        self.store.load_state_dict(init_tf_params(shapes, seed=self.seed))
        self.pipeline = EpisodePipeline(self.cfg, self.store, emb, lr=self.learning_rate0,
                                        decay_steps=self.decay_steps, decay_rate=self.decay_rate, seed=self.seed)
        self.word_embedding = emb if emb is not None else self.store.view("word_embedding")
        self.pos1_embedding = self.store.view("pos1_embedding")
        self.pos2_embedding = self.store.view("pos2_embedding")
        self._trainer = None

    def feed(self, inputs, query_label, keep_prob=None):
        """The ``feed_dict`` of model/base.py:145-151."""
        self.pipeline.set_inputs(inputs, query_label)
        self._keep = self.keepProb if keep_prob is None else keep_prob
        self.input_words, self.input_pos1 = self.pipeline.word, self.pipeline.pos1
        self.input_pos2, self.mask_padding = self.pipeline.pos2, self.pipeline.mask
        self.query_label = self.pipeline.label

    def get_embedding(self):
        """[X, L, E + 2*5] embedded batch (model/base.py:76-86) - valid after ``forward``."""
        c = self.cfg
        return self.pipeline.x.view(c.X, c.L, c.D)

    def forward(self):
        raise NotImplementedError

    def build_loss(self):
        raise NotImplementedError

    def build_optimize(self):
        """Backward + staircase-decayed TF Adam (model/base.py:94-99) on the batch just run forward."""
        self.pipeline.backward()
        self.pipeline.adam()
        self.learning_rate = float(self.store.hyper[1].item())
        self.global_step = int(self.store.global_step.item())

    def build_predict(self):
        self.predict_test = self.probs.argmax(dim=-1)        # first max wins, like tf.argmax
        self.predict = torch.round(self.probs)               # half-to-even, like tf.round

    def build_accuracy(self):
        self.accuracy = float(self.pipeline.out2[1].item())  # fused into the NTN head kernel

    def build_summary(self):
        self.summary = {"loss": self.loss, "accuracy": self.accuracy, "learning_rate": self.learning_rate}

    # -------------------------------------------------------------------------------- one step
    def run_step(self, inputs, query_label, train=True):
        """What one ``sess.run([loss, accuracy, optimize, ...])`` does (model/base.py:143-152)."""
        self.feed(inputs, query_label, self.keepProb if train else 1.0)
        if train:
            self.pipeline.cfg.keep_prob = float(self._keep)     # eval runs forward(train=False): keep = 1 (base.py:199)
        self._train_mode = train
        self.forward()
        self.build_predict()
        self.build_accuracy()
        self.build_loss()
        if train:
            self.build_optimize()
        self.build_summary()
        return self.loss, self.accuracy

    def _sample(self, loader):
        c = self.cfg
        return loader.next_batch(c.B, c.N, c.K, c.Q)

    # -------------------------------------------------------------------------------- loops
    def _log_scalars(self, step, loss, accuracy):
        """The three tf.summary scalars of model/base.py:110-113 as one JSON line per logging point."""
        if self.scalar_log:
            import json
            with open(self.scalar_log, "a") as f:
                f.write(json.dumps({"step": int(step), "loss": float(loss), "accuracy": float(accuracy),
                                    "learning_rate": float(self.learning_rate)}) + "\n")

    def make_trainer(self, train_data_loader):
        """EpisodeTrainer over this model's pipeline and the loader's device sampler; data parallel when the
        process was launched with WORLD_SIZE > 1 (one process per GPU, torchrun)."""
        from .. import dist as ddist
        from ..trainer import EpisodeTrainer
        rank, _local, world = ddist.init_from_env()
        smp = train_data_loader._device()
        rows = self.pipeline.word_table.shape[0] if self.pipeline.word_table is not None else int(self.vocab_size)
        smp.validate_ids(int(rows), self.cfg.L)      # out-of-range ids would be out-of-bounds reads in the embedding kernels
        smp.load_host_streams()
        self.pipeline.cfg.keep_prob = float(self.keepProb)
        if world > 1:                       # every rank starts from rank 0's weights
            ddist.broadcast_(self.store.params)
            self.store.refresh_bf16()
        self._trainer = EpisodeTrainer(self.pipeline, smp, rank=rank, world=world, cuda_graph=True)
        return self._trainer

    def train(self, dataloader, model_dir_path, model_name="inductionNetwork",
              train_iter=30000, val_iter=1000, val_step=2000, test_iter=3000):
        train_data_loader, val_data_loader = dataloader
        if not exists(model_dir_path):
            makedirs(model_dir_path)
        curr_iter, best_acc = 0, 0.0
        from ..dist import env_world
        rank, _, world = env_world()
        fast = self.use_cuda_graph or world > 1
        trainer = self.make_trainer(train_data_loader) if fast else None
        say = print if rank == 0 else (lambda *a, **k: None)
        say("training start ..")
        iter_loss, iter_right, iter_sample = 0.0, 0.0, 0.0
        for it in range(curr_iter, curr_iter + train_iter):
            if fast:
                # sampler + fwd + bwd + (all-reduce) + Adam as one graph replay; the running means of
                # model/base.py:156-158 live on the device and are read only where the reference prints them
                trainer.step()
                if it % 100 == 0 or it % val_step == 0 or (it + 1) % val_step == 0:
                    mean_loss, mean_acc, n = trainer.read_running(reset=(it % val_step == 0))
                    self.loss, self.accuracy = trainer.read()
                    self.learning_rate = float(self.store.hyper[1].item())
                    self.global_step = int(self.store.global_step.item())
                    iter_loss, iter_right, iter_sample = mean_loss * n, mean_acc * n, float(max(n, 1))
            else:
                inputs, query_label = self._sample(train_data_loader)
                curr_loss, curr_acc = self.run_step(inputs, query_label, train=True)
                iter_loss += curr_loss
                iter_right += curr_acc
                iter_sample += 1
            if it % 100 == 0:
                say('[train] step:{0:4} | loss: {1:2.6f}, accuracy: {2:3.2f}%'.format(
                    it + 1, iter_loss / iter_sample, 100 * iter_right / iter_sample) + '\r')
                if rank == 0:
                    self._log_scalars(it + 1, iter_loss / iter_sample, iter_right / iter_sample)
            if it % val_step == 0:
                iter_loss, iter_right, iter_sample = 0.0, 0.0, 0.0
            if (it + 1) % val_step == 0:
                say("evaluate start ..")
                if fast:
                    trainer.suspend()      # host streams back where the reference's loop would have them
                acc_val = self.eval(val_data_loader, None, val_iter)
                if acc_val > best_acc:
                    say("Best checkpoint.[EVAL] accuracy :{}".format(acc_val))
                    ckpt_dir = join(model_dir_path, "checkpoint")
                    if not exists(ckpt_dir):
                        makedirs(ckpt_dir)
                    if rank == 0:
                        self.global_step = int(self.store.global_step.item())
                        self.save(join(ckpt_dir, "%s-%d.npz" % (model_name, self.global_step)))
                    best_acc = acc_val
        if fast:
            trainer.suspend()
        say("\n####################\n")
        say("Finish training :" + model_name)
        test_acc = self.eval(val_data_loader, None, test_iter)
        say("Test accuracy: {}".format(test_acc))
        if fast:
            trainer.close()
            self._trainer = None
        return test_acc

    def eval(self, val_data_loader, sess, val_iter):
        """``sess`` is kept as an ignored positional (model/base.py:186).  The reference forgets to feed
        ``mask_padding`` here (:195-199) and cannot run; this feeds the mask."""
        iter_right_val, iter_sample_val = 0.0, 0.0
        for it_val in range(val_iter):
            inputs_val, query_label_val = self._sample(val_data_loader)
            _, curr_acc_val = self.run_step(inputs_val, query_label_val, train=False)
            iter_right_val += curr_acc_val
            iter_sample_val += 1
            if it_val % 100 == 0 and self._quiet_rank() == 0:
                print('[EVAL] step: {0:4} | accuracy: {1:3.2f}%'.format(
                    it_val + 1, 100 * iter_right_val / iter_sample_val) + '\r')
        return iter_right_val / iter_sample_val

    @staticmethod
    def _quiet_rank():
        from ..dist import env_world
        return env_world()[0]

    # -------------------------------------------------------------------------------- checkpoint
    def save(self, path):
        """Checkpoint keyed by the TF variable names + Adam slots + global_step."""
        sd = {k: v.cpu().numpy() for k, v in self.store.state_dict().items()}
        sd.update({"adam_m/" + k: v.cpu().numpy() for k, v in self.store._export(self.store.m).items()})
        sd.update({"adam_v/" + k: v.cpu().numpy() for k, v in self.store._export(self.store.v).items()})
        sd["global_step"] = np.int64(self.store.global_step.item())
        if self.embed is not None:
            # the frozen word table is not a trainable and is not stored (the reference's Saver stores it): its shape
            # and a CRC travel with the checkpoint so that restore() can refuse a different embedding file
            sd["pred_embed/shape"] = np.asarray(np.asarray(self.embed).shape, dtype=np.int64)
            sd["pred_embed/crc32"] = np.int64(self._embed_crc())
        np.savez(path, **sd)

    def _embed_crc(self):
        import zlib
        return zlib.crc32(np.ascontiguousarray(np.asarray(self.embed), dtype=np.float32).tobytes()) & 0xFFFFFFFF

    def restore(self, path):
        z = np.load(path)
        if "pred_embed/crc32" in z and self.embed is not None:
            shape = tuple(int(x) for x in z["pred_embed/shape"])
            if shape != tuple(np.asarray(self.embed).shape) or int(z["pred_embed/crc32"]) != self._embed_crc():
                raise ValueError("checkpoint %s was trained with a different word embedding table (shape %s)" % (path, shape))
        self.store.load_state_dict({k: torch.from_numpy(z[k]) for k in self.store.tf_shapes})
        for slot, buf in (("adam_m/", self.store.m), ("adam_v/", self.store.v)):
            if slot + "M" in z:
                for k in self.store.tf_shapes:
                    t = torch.from_numpy(z[slot + k])
                    t = t.permute(0, 2, 1).contiguous() if k == "M" else t
                    self.store.view(k, buf).copy_(t.to(self.store.device))
        self.store.global_step.fill_(int(z["global_step"]))
        self.global_step = int(z["global_step"])
