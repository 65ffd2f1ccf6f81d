"""Job runner — mirror of the part of lib/main.js the stages depend on (lib/main.js:99-170), and of js/lib/main.js,
which cannot be executed on this image.  The control plane around it (AMQP, Prometheus, `/health`, telemetry
transport) is out of scope; queue, telemetry, S3 client and codec are injected, as in the JS file.

Kept from the reference, in its order:
  * stages are created by name, in order, from factories `(config, emitter, logger) -> stage`, and a factory that does
    not return a callable is an error (lib/main.js:99-115);
  * a job whose `<id>/original/done` object already exists in `triton-staging` skips every stage and only publishes
    the convert job again (lib/main.js:117-124,153-155) — that is what the upload stage writes `done` for;
  * each stage receives the job with the previous stage's result as `lastStage` (lib/main.js:129-140);
  * an error whose code is 'ERRDLSTALL' acks (drops) the message; any other error emits status ERRORED (6) and
    nacks (lib/main.js:141-151);
  * a failed publish of `v1.convert` is logged and the message is left neither acked nor nacked (lib/main.js:157-166).
"""
from __future__ import annotations

import os
from datetime import datetime, timezone

from . import Job, TelemetryStatus
from . import download, process, upload

STAGES = ("download", "process", "upload")
BUCKET = "triton-staging"
_FACTORIES = {"download": download.factory, "process": process.factory, "upload": upload.factory}


async def factory(config, logger, factories=None):
    """-> async handle(message).  message: .content (the decoded job dict is config['codec'].decode_download(content)),
    .ack(), .nack().  config: 'queue' (publish(topic, bytes)), 'telem' (emitStatus / emitProgress), 's3' (getObject),
    'codec' (decode_download / encode_convert), plus what the stage factories read."""
    factories = factories or _FACTORIES
    telem = config["telem"]

    async def handle(message):
        job = config["codec"].decode_download(message.content)
        job_id = job["media"]["id"]
        await telem.emitStatus(job_id, TelemetryStatus.DOWNLOADING)
        try:
            table = {}
            for name in STAGES:
                fn = await factories[name](config, {}, logger)
                if not callable(fn):
                    raise TypeError(f"Invalid stage '{name}' return value was not a function")
                table[name] = fn
            staged = True
            try:
                logger.info("checking s3 bucket to see if files already exist for id", job_id)
                config["s3"].getObject(BUCKET, os.path.join(job_id, "original/", "done"))
            except Exception as err:  # noqa: BLE001 — any failure means "not staged yet", as in the reference
                logger.info("failed to find done file in staging", str(err))
                staged = False
            if not staged:
                last = {}
                for name in STAGES:
                    logger.info(f"invoking stage '{name}'")
                    last = await table[name](Job(job, lastStage=last))
            else:
                logger.warn("skipping download due to files existing in triton-staging")
        except Exception as err:  # noqa: BLE001
            logger.error("failed to invoke stage:", str(err))
            if getattr(err, "code", None) == "ERRDLSTALL":
                return message.ack()
            await telem.emitStatus(job_id, TelemetryStatus.ERRORED)
            return message.nack()
        payload = {"createdAt": datetime.now(timezone.utc).isoformat(), "media": job["media"]}
        try:
            config["queue"].publish("v1.convert", config["codec"].encode_convert(payload))
        except Exception as err:  # noqa: BLE001
            return logger.error("failed to create job:", str(err))
        message.ack()

    return handle
