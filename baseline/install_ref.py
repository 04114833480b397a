"""Copy the reference's torch env -- UNMODIFIED -- into ``baseline/_ref`` so that the GPU box (where /root/reference
does not exist) can time it beside the CUDA env (baseline B3).

Files (all of /root/reference): torch_impl/env/{__init__,env,wrappers}.py, common/{__init__,constants}.py and the five
bundled agent checkpoints sample_models/dqn-agent-{1..5}.safetensors (the evaluator's opponents).
``baseline/_ref`` is git-ignored (reference sources never enter this repo's history) but travels with gpurun.
Run by ``__graft_entry__.build()`` whenever /root/reference is present; a no-op otherwise.
"""
import hashlib
import json
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
REF = '/root/reference'
DEST = os.path.join(HERE, '_ref')
FILES = ['torch_impl/__init__.py', 'torch_impl/env/__init__.py', 'torch_impl/env/env.py', 'torch_impl/env/wrappers.py',
         'common/__init__.py', 'common/constants.py'] + [
    # the five bundled baseline agents of drone_evaluator.py:31-37 (weights, 168 KB): opponents of the evaluator loop
    f'sample_models/dqn-agent-{i}.safetensors' for i in range(1, 6)]


def install(ref=REF, dest=DEST):
    """Returns the manifest {relative path: sha256} of what was copied, or None when the reference is absent."""
    if not os.path.isdir(ref):
        return None
    manifest = {}
    for rel in FILES:
        src = os.path.join(ref, rel)
        dst = os.path.join(dest, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if os.path.exists(src):
            shutil.copyfile(src, dst)
            manifest[rel] = hashlib.sha256(open(src, 'rb').read()).hexdigest()
        elif rel.endswith('__init__.py'):
            open(dst, 'w').close()          # namespace marker the reference does not ship
            manifest[rel] = 'created-empty'
    with open(os.path.join(dest, 'MANIFEST.json'), 'w') as f:
        json.dump({'source': ref, 'files': manifest}, f, indent=1)
    return manifest


def available(dest=DEST):
    return os.path.exists(os.path.join(dest, 'torch_impl', 'env', 'env.py'))


if __name__ == '__main__':
    print(json.dumps(install(), indent=1))
