"""Host-side mirror of cvpce/production.py for the hot path: ProposalGenerator and Classifier with the reference's
constructor arguments, method names and return conventions; the per-box / per-batch python loops are replaced by
single launches of the sm_100a kernels.  PlanogramComparator / PlanogramEvaluator (CPU graph matching + RANSAC) are
out of scope and not mirrored; the reference's own classes accept these objects unchanged.
"""
import torch
from torch.utils.data import DataLoader

from . import classification, datautils


def scale_to_tanh(tensor):
    """cvpce/utils.py:280-281"""
    return tensor * 2 - 1


def _gp_annotated_collate(samples):
    """datautils.py:297-299 gp_annotated_collate_fn: (images stacked, images list, categories list, annotations list)."""
    imgs, raw, cats, anns = zip(*samples)
    return torch.stack(imgs), list(raw), list(cats), list(anns)


class ProposalGenerator:
    """production.py:8-20"""

    def __init__(self, detector, device=torch.device('cuda'), confidence_threshold=0.5):
        self.detector = detector
        self.device = device
        self.condfidence_threshold = confidence_threshold  # (sic) attribute name of the reference

    def generate_proposals(self, image):
        res = self.detector(image[None].to(self.device))[0]
        return res['boxes'][res['scores'] > self.condfidence_threshold]

    def generate_proposals_and_images(self, image):
        """Returns (boxes f32[K,4] on the detector device, crops f32[K,3,256,256]).  The crops stay on the device
        (the reference builds them on the CPU and Classifier.classify uploads them again)."""
        boxes = self.generate_proposals(image)
        S = datautils.CLASSIFICATION_IMAGE_SIZE
        if not len(boxes):
            return boxes, torch.empty((0, 3, S, S), device=self.device)
        return boxes, datautils.crops_for_classification(image.to(self.device), boxes)


class Classifier:
    """production.py:22-74"""

    def __init__(self, encoder, sample_set, device=torch.device('cuda'), emb_device=torch.device('cuda'), batch_size=32,
                 num_workers=8, k=1, load=None, verbose=False):
        self.batch_size = batch_size
        self.num_workers = num_workers
        self.device = device
        self.emb_device = emb_device
        self.k = k
        self.encoder = encoder
        if load is None:
            self.embedding, self.annotations = self.build_index(sample_set, verbose)
        else:
            self.embedding, self.annotations = self.load_index(load)
        self._prepared = None
        self._sharded = None

    def build_index(self, sample_set, verbose=False):
        loader = DataLoader(sample_set, batch_size=self.batch_size, num_workers=self.num_workers,
                            collate_fn=_gp_annotated_collate, pin_memory=True)
        chunks, annotations = [], []
        with torch.no_grad():
            for i, (imgs, _, _, anns) in enumerate(loader):
                if verbose and i % 100 == 0:
                    print(i)
                chunks.append(self.encoder(imgs.to(device=self.device)).detach().to(device=self.emb_device))
                annotations += anns
        if chunks:
            embedding = torch.cat(chunks)  # one concatenation instead of the reference's O(G^2) grow-by-append
        else:
            embedding = torch.empty((0, self.encoder.embedding_size), dtype=torch.float, device=self.emb_device)
        return embedding, annotations

    def save_index(self, pth):
        torch.save({'embedding': self.embedding, 'annotations': self.annotations}, pth)

    def load_index(self, pth):
        idx = torch.load(pth, map_location=self.emb_device)
        return idx['embedding'], idx['annotations']

    def shard_gallery(self, world, rank, group=None, exchange="nccl"):
        """Multi-GPU gallery: keep only this rank's row range of the index on the device; classify() then
        searches the shard, exchanges the per-rank top-k blocks and merges, and every rank maps the same GLOBAL indices to
        annotations (the annotation list stays whole on every rank).  Returns the ShardedGallery."""
        from . import dist as cdist
        G = int(self.embedding.shape[0])
        lo, hi = cdist.shard_range(G, world, rank)
        self._sharded = cdist.ShardedGallery(self.embedding[lo:hi].to(self.emb_device), lo, world, rank, group, exchange=exchange,
                                             global_rows=G)
        return self._sharded

    def prepared_gallery(self):
        if self._prepared is None or self._prepared.f32.data_ptr() != self.embedding.data_ptr():
            self._prepared = classification.PreparedGallery(self.embedding.to(self.emb_device))
        return self._prepared

    def embed(self, images):
        """encoder over all images in trunk batches of batch_size -> f32[K, embedding_size] on emb_device."""
        out = torch.empty((len(images), self.encoder.embedding_size), dtype=torch.float, device=self.emb_device)
        with torch.no_grad():
            for i in range(0, len(images), self.batch_size):
                batch = scale_to_tanh(images[i:i + self.batch_size].to(device=self.device))
                out[i:i + self.batch_size] = self.encoder(batch).detach().to(device=self.emb_device)
        return out

    def classify(self, images, return_embedding=False):
        """-> list (len K) of lists (len k) of annotation strings [, embeddings f32[K,D]].  One nearest-neighbour launch
        and one device->host copy for the whole call (the reference syncs once per element, production.py:70)."""
        if len(images) == 0:
            res = []
            emb = torch.empty((0, self.encoder.embedding_size), dtype=torch.float, device=self.emb_device)
            return (res, emb) if return_embedding else res
        emb = self.embed(images)
        if getattr(self, "_sharded", None) is not None:
            # every rank must pass the SAME images here (the shards are searched for all queries and the per-rank blocks merged);
            # image data-parallel ranks keep the whole gallery on every GPU instead (dist.partition_images)
            _, nearest = self._sharded.search(emb, self.k)
        else:
            _, nearest = classification.similarity_topk(self.embedding, emb, self.k, prepared=self.prepared_gallery())
        nearest = nearest.cpu()
        if bool((nearest < 0).any()):  # padding (-1) can only survive when the gallery holds fewer rows than requested
            raise RuntimeError("Classifier.classify: the search returned padding indices (gallery smaller than k?)")
        nearest = nearest.tolist()
        res = [[self.annotations[j] for j in row] for row in nearest]
        if return_embedding:
            return res, emb
        return res
