#!/usr/bin/env python
"""Run one of the fork's own scripts UNMODIFIED against this package.

    python tools/run_fork_script.py examples/gcn.py
    python tools/run_fork_script.py examples/ConvexPruning.py -- --dataset Cora \
        --num_pre_epochs 1 --num_epochs 2 --PruningTimes 1 --NumLayers 2
    python tools/run_fork_script.py ConvexPruning.py -- --dataset Cora --modelName GAT ...

The script file is executed as is (``runpy.run_path(..., run_name='__main__')``) from the
read-only reference tree — ``/root/reference`` in the build container; on a GPU box the same
files are unpacked from ``baseline/_ref/fork_scripts.tar.gz`` (written by
``__graft_entry__.build()``, git-ignored, never part of the repo's history).  Nothing in the
script is edited; what this launcher supplies is the ENVIRONMENT the fork assumed:

* ``torch_geometric`` = this package (first on ``sys.path``);
* inert stand-ins for third-party modules that are imported at module level but not
  installed here and not used on the GCN / GAT path: ``pyts.image``, ``cupy``,
  ``cupy.core.dlpack``, ``torch_cluster``, ``community``, ``pathos``, ``pyamg`` and
  ``matplotlib`` (any attribute resolves; calling one raises);
* a scratch working directory with the result folders the scripts write into
  (``Results/<dataset>Convergence``, ``checkpoint``) — their dataset roots
  (``/git/data/GraphData/<name>``, ``../data/<name>``) do not exist, so the dataset classes
  fall back to the seeded synthetic stand-ins and warn about it;
* ``TORCH_FORCE_NO_WEIGHTS_ONLY_LOAD=1``: the scripts ``torch.load`` whole pickled modules,
  which torch >= 2.6 refuses by default (torch version drift, not part of the hot path).
"""
import argparse
import os
import runpy
import sys
import tarfile
import tempfile
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, 'pytorch_geometric-1_b200')
REFERENCE = '/root/reference'
ARCHIVE = os.path.join(ROOT, 'baseline', '_ref', 'fork_scripts.tar.gz')
STUBS = ['pyts', 'pyts.image', 'cupy', 'cupy.core', 'cupy.core.dlpack', 'torch_cluster',
         'community', 'pathos', 'pathos.multiprocessing', 'pyamg', 'matplotlib',
         'matplotlib.pyplot', 'matplotlib.ticker', 'matplotlib.cm', 'matplotlib.colors']


class _Inert(types.ModuleType):
    """A module in which every name resolves to a callable that raises when used."""

    def __getattr__(self, name):
        if name.startswith('__'):
            raise AttributeError(name)
        full = self.__name__ + '.' + name

        def _missing(*args, **kwargs):
            raise ImportError('{} is a stand-in: the module is not installed in this image and is '
                              'not on the message-passing path'.format(full))
        _missing.__name__ = name
        return _missing


def install_stubs():
    import importlib
    added = []
    for name in STUBS:
        top = name.split('.')[0]
        if top not in sys.modules:
            try:
                importlib.import_module(top)
            except ImportError:
                pass
        if top in sys.modules and not isinstance(sys.modules[top], _Inert):
            continue                                    # the real package exists: leave it alone
        if name not in sys.modules:
            mod = _Inert(name)
            mod.__path__ = []                           # lets `import a.b` treat it as a package
            sys.modules[name] = mod
            added.append(name)
            if '.' in name:
                parent, leaf = name.rsplit('.', 1)
                setattr(sys.modules[parent], leaf, mod)
    return added


def reference_root():
    """The fork's tree: the read-only mount, else the archive unpacked into a temp dir."""
    if os.path.isdir(REFERENCE):
        return REFERENCE
    if os.path.exists(ARCHIVE):
        dst = tempfile.mkdtemp(prefix='fork_scripts_')
        with tarfile.open(ARCHIVE) as tf:
            tf.extractall(dst)
        return dst
    return None


def pack_reference_scripts():
    """Called by ``__graft_entry__.build()`` where /root/reference exists: archive the fork's
    top-level and ``examples/`` python files so that the GPU box can execute them."""
    if not os.path.isdir(REFERENCE):
        return False
    os.makedirs(os.path.dirname(ARCHIVE), exist_ok=True)
    with tarfile.open(ARCHIVE, 'w:gz') as tf:
        for sub in ('', 'examples'):
            folder = os.path.join(REFERENCE, sub)
            for fn in sorted(os.listdir(folder)):
                if fn.endswith('.py'):
                    tf.add(os.path.join(folder, fn), arcname=os.path.join(sub, fn))
    return True


def run(script, argv=(), workdir=None):
    root = reference_root()
    if root is None:
        raise FileNotFoundError('neither {} nor {} exists'.format(REFERENCE, ARCHIVE))
    path = os.path.join(root, script)
    if not os.path.exists(path):
        raise FileNotFoundError(path)
    os.environ.setdefault('TORCH_FORCE_NO_WEIGHTS_ONLY_LOAD', '1')
    for p in (os.path.dirname(path), root, PKG):        # PKG ends up first
        if p in sys.path:
            sys.path.remove(p)
        sys.path.insert(0, p)
    stubs = install_stubs()
    workdir = workdir or tempfile.mkdtemp(prefix='fork_run_')
    for ds in ('Cora', 'Citeseer', 'Pubmed', 'CoraFull', 'ENZYMES', 'MUTAG', 'PPI', 'Reddit',
               'Amazon', 'MNIST'):
        os.makedirs(os.path.join(workdir, 'Results', ds + 'Convergence'), exist_ok=True)
    os.makedirs(os.path.join(workdir, 'Results', 'PartitionResults'), exist_ok=True)
    os.makedirs(os.path.join(workdir, 'checkpoint'), exist_ok=True)
    old_cwd, old_argv = os.getcwd(), sys.argv
    os.chdir(workdir)
    sys.argv = [path] + list(argv)
    print('[run_fork_script] {} {} (cwd {}; stand-in modules: {})'.format(
        path, ' '.join(argv), workdir, ', '.join(stubs) or 'none'), flush=True)
    try:
        return runpy.run_path(path, run_name='__main__')
    finally:
        os.chdir(old_cwd)
        sys.argv = old_argv


def main():
    # hand-parsed so that everything after `--` reaches the script untouched
    argv = sys.argv[1:]
    rest = []
    if '--' in argv:
        k = argv.index('--')
        argv, rest = argv[:k], argv[k + 1:]
    ap = argparse.ArgumentParser(description=__doc__.split('\n')[0])
    ap.add_argument('script', help='path relative to the fork root, e.g. examples/gcn.py')
    ap.add_argument('--workdir', default=None)
    args = ap.parse_args(argv)
    run(args.script, rest, args.workdir)
    print('[run_fork_script] finished OK')


if __name__ == '__main__':
    main()
